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

# ---- full capture -> metrics + traffic -----------------------------------------------------------------------------------------------
rep = os.path.join(OUT, "ncu_full_gemv.ncu-rep")
if os.path.exists(rep):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr = rows[0]; units = rows[1]; body = rows[2:]
    want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
            "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size", "smsp__inst_executed.sum",
            "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
            "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
            "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct"]
    idx = {h: i for i, h in enumerate(hdr)}
    kn = idx.get("Kernel Name")
    tot_dram = 0.0
    with open(dst("ncu_full_gemv.txt"), "w") as o:
        o.write("# ncu --set full --clock-control none --import-source on -k regex:'k_gemv_q4k_mma|k_gemv_stream' -s 20 -c 6 -- python bench.py --steps 2 --warmup 3 --no-pp --no-cpu --no-graph\n")
        o.write("# six consecutive decode MUL_MAT launches of Llama-3-8B Q4_K_M (layers 4-5: wq+wk / wv or wq+wk+wv, wo, gate+up, down, ...); serialised, cold-cache, no PDL overlap\n\n")
        for r in body:
            o.write(f"{r[kn]}\n")
            for w in want:
                if w in idx:
                    o.write(f"    {w:<90s} {r[idx[w]]} {units[idx[w]]}\n")
            o.write("\n")
            for w in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                v = float(r[idx[w]].replace(",", "")); u = units[idx[w]].lower()
                tot_dram += v * (1e6 if u.startswith("mbyte") else 1e3 if u.startswith("kbyte") else 1e9 if u.startswith("gbyte") else 1)
    print("wrote", dst("ncu_full_gemv.txt"), "dram bytes captured", tot_dram)
    json.dump({"source": f"profiles/{tag}_ncu_full_gemv.txt", "dram_bytes_captured": tot_dram, "algorithmic_bytes_captured": 146628608, "traffic_over_algorithmic": round(tot_dram / 146628608, 4),
               "note": "algorithmic bytes: the q4_K / q6_K blocks of the six captured matrices (wq+wk+wv 14.2 MB, wo 9.4, gate+up 66.1, down 33.0, wq+wk+wv, wo) of layers 4-5"},
              open(dst("traffic.json"), "w"), indent=1)
