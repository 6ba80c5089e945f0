#!/usr/bin/env python
"""Turn what scripts/gpu_round.sh left in gpurun_out/ into the tracked profiles/<tag>_* files: bench lines, pytest log, the ncu launch
list (raw + per-kernel shares), the metrics of the full capture (ncu -i ... --page raw --csv, read here) with DRAM traffic against the
algorithmic bytes, llama-bench tables.  usage: collect_profiles.py r1_run7"""
import csv, collections, json, os, re, shutil, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT, PROF = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")
tag = sys.argv[1]
dst = lambda name: os.path.join(PROF, f"{tag}_{name}")


def copy(src, name):
    p = os.path.join(OUT, src)
    if os.path.exists(p):
        shutil.copy(p, dst(name)); print("wrote", dst(name))


for src, name in (("bench.json", "bench.json"), ("bench_reference.json", "bench_reference.json"), ("pytest_gpu.log", "pytest_gpu.log"), ("smoke.log", "smoke.log"),
                  ("gpu.txt", "gpu.txt"), ("cpu.txt", "cpu.txt"), ("ncu_launches.csv", "ncu_launches.csv")):
    copy(src, name)
for f in sorted(os.listdir(OUT)):
    if f.startswith("llama_bench_") and f.endswith(".txt"):
        copy(f, f)
    m = re.match(r"test-backend-ops_test_(\w+?)(_glue|_repack)?\.log", f)
    if m:                                                                   # one summary line per op instead of thousands of case lines
        txt = re.sub(r"\x1b\[[0-9;]*m", "", open(os.path.join(OUT, f)).read())
        ok = len(re.findall(r": OK$", txt, re.M)); ns = txt.count("not supported"); fail = len(re.findall(r"FAIL", txt))
        tail = [l for l in txt.splitlines() if "tests passed" in l or "Backend" in l][-2:]
        with open(dst("test-backend-ops.summary.txt"), "a") as o:
            o.write(f"{m.group(1)}{m.group(2) or ''}: {ok} OK, {fail} FAIL, {ns} not supported by the plugin (run nowhere); {' | '.join(tail)}\n")

# ---- launch list -> shares --------------------------------------------------------------------------------------------------------
p = os.path.join(OUT, "ncu_launches.csv")
if os.path.exists(p):
    lines = [l for l in open(p) if not l.startswith("==")]
    agg = collections.OrderedDict()
    for r in csv.DictReader(lines):
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(r["Metric Value"].replace(",", "")); u = r["Metric Unit"]
        v = v / 1e3 if u == "ns" else v * 1e3 if u == "ms" else v
        k = re.sub(r"\(.*", "", r["Kernel Name"])
        a = agg.setdefault(k, [0, 0.0]); a[0] += 1; a[1] += v
    tot = sum(a[1] for a in agg.values())
    with open(dst("ncu_launch_summary.csv"), "w") as o:
        o.write("# ncu --metrics gpu__time_duration.sum --clock-control none -c 700, python bench.py --steps 2 --warmup 3 --no-pp --no-cpu --no-graph (first 700 launches)\n")
        o.write("# per-launch times are cold-cache and serialised: shares, not absolutes\nkernel,launches,total_us,avg_us,share\n")
        for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            o.write(f"{k},{n},{t:.1f},{t / n:.2f},{t / tot:.3f}\n")
    print("wrote", dst("ncu_launch_summary.csv"))

# ---- full captures -> metrics (+ DRAM traffic against the algorithmic bytes for the dominant kernel) --------------------------------------------
KEEP = re.compile(r"^(gpu__time_duration\.sum|dram__bytes_(read|write)\.sum|dram__throughput\.avg\.pct_of_peak_sustained_elapsed|sm__throughput\.avg\.pct_of_peak_sustained_elapsed|"
                  r"sm__warps_active\.avg\.pct_of_peak_sustained_active|smsp__issue_active\.avg\.pct_of_peak_sustained_active|smsp__inst_executed\.sum|launch__(registers_per_thread|grid_size|block_size|"
                  r"shared_mem_per_block_dynamic|occupancy_limit_\w+|waves_per_multiprocessor)|sm__pipe_tensor\w*\.avg\.pct_of_peak_sustained_active|sm__inst_executed_pipe_(tensor\w*|uniform|lsu|alu|fma\w*|xu)\.sum|"
                  r"sm__pipe_(alu|fma|fmaheavy|shared)_cycles_active\.avg\.pct_of_peak_sustained_active|lts__throughput\.avg\.pct_of_peak_sustained_elapsed|lts__t_bytes\.sum|lts__t_sector_hit_rate\.pct|"
                  r"l1tex__throughput\.avg\.pct_of_peak_sustained_(active|elapsed)|l1tex__data_pipe_lsu_wavefronts_mem_shared\.sum|l1tex__data_bank_conflicts_pipe_lsu_mem_shared\.sum|"
                  r"smsp__average_warps_issue_stalled_\w+_per_issue_active\.ratio|smsp__cycles_active\.avg|sm__cycles_elapsed\.max)$")


def to_bytes(v, u):
    v = float(v.replace(",", "")); u = u.lower()
    return v * (1e6 if u.startswith("mbyte") else 1e3 if u.startswith("kbyte") else 1e9 if u.startswith("gbyte") else 1)


def summarize(rep_name, out_name, header):
    rep = os.path.join(OUT, rep_name)
    if not os.path.exists(rep):
        return None
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units, body = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(hdr)}
    kn = idx.get("Kernel Name")
    per_launch = []
    with open(dst(out_name), "w") as o:
        o.write(header)
        for r in body:
            o.write(f"{r[kn]}\n")
            for h in hdr:
                if KEEP.match(h) and r[idx[h]] not in ("", "n/a"):
                    if "issue_stalled" in h and float(r[idx[h]].replace(",", "") or 0) < 0.05:
                        continue                                                  # only the stall reasons that matter
                    o.write(f"    {h:<100s} {r[idx[h]]} {units[idx[h]]}\n")
            o.write("\n")
            per_launch.append((r[kn], sum(to_bytes(r[idx[w]], units[idx[w]]) for w in ("dram__bytes_read.sum", "dram__bytes_write.sum"))))
    print("wrote", dst(out_name))
    return per_launch


common = "ncu --set full --clock-control none --import-source on, python bench.py --steps 2 --warmup 3 --no-cpu --no-graph --no-llama-bench (serialised, cold-cache replays: metrics, not times)"
chain = summarize("ncu_full_chain.ncu-rep", "ncu_full_decode_chain.txt",
                  f"# {common} --no-pp -k regex:k_decode_chain -s 4 -c 2\n# the persistent decode chain of layers 4 and 5 of Llama-3-8B Q4_K_M (wo, +residual, ffn_norm, gate/up, SwiGLU, down, +ffn_inp, attn_norm, next wq/wk/wv)\n\n")
if chain:
    sys.path.insert(0, ROOT)
    import bench as Bn
    cfg = Bn.MODELS["8b"]; types, _ = Bn.q4_k_m_layer_types(cfg, "q4_k_m"); shapes = Bn.weight_shapes(cfg)
    alg = []
    for il in (4, 5):                                                             # chain il: wo / gate / up / down of layer il, wq / wk / wv of layer il + 1
        alg.append(sum(Bn.wbytes(types[il][n], *shapes[n]) for n in ("wo", "gate", "up", "down")) + sum(Bn.wbytes(types[il + 1][n], *shapes[n]) for n in ("wq", "wk", "wv")))
    tot = sum(b for _, b in chain)
    json.dump({"source": f"profiles/{tag}_ncu_full_decode_chain.txt", "kernel": "k_decode_chain", "launches_captured": len(chain), "dram_bytes_captured": tot,
               "algorithmic_bytes_captured": sum(alg[:len(chain)]), "traffic_over_algorithmic": round(tot / sum(alg[:len(chain)]), 4),
               "note": "algorithmic bytes: the q4_K / q6_K blocks of the seven matrices each captured chain launch streams (layers 4 and 5); activations and hand-over traffic are < 0.1 %"},
              open(dst("traffic.json"), "w"), indent=1)
    print("wrote", dst("traffic.json"))
summarize("ncu_full_gemm.ncu-rep", "ncu_full_gemm.txt", f"# {common} -k regex:k_gemm_q -s 7 -c 8\n# prefill MUL_MAT (tcgen05 GEMM with in-kernel dequantization) of a 512-token ubatch, Llama-3-8B Q4_K_M: q4_K (<12>) and q6_K (<14>) launches\n\n")
summarize("ncu_full_fattn_tc.ncu-rep", "ncu_full_fattn_tc.txt", f"# {common} -k regex:k_fattn_tc -s 2 -c 2\n# prefill FLASH_ATTN_EXT (tcgen05), 512 queries x 512 keys, 32:8 heads of 128\n\n")
summarize("ncu_full_fattn_gqa.ncu-rep", "ncu_full_fattn_gqa.txt", f"# {common} --no-pp --depth 8192 -k regex:k_fattn_gqa -s 40 -c 2\n# decode FLASH_ATTN_EXT over 8448 keys (GQA-grouped mma.sync kernel, TMA tiles)\n\n")
summarize("ncu_full_gemv.ncu-rep", "ncu_full_gemv.txt", f"# {common} --no-pp with B200_BENCH_NO_CHAIN=1 -k regex:'k_gemv_q4k_mma|k_gemv_stream' -s 20 -c 6\n\n")
for f in ("umma_peak.txt", "sanitize_memcheck.txt", "sanitize_racecheck.txt", "chain_bench.log"):
    copy(f, f)
