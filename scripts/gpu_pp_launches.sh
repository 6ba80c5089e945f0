#!/bin/bash
# one gpurun call: ncu launch list (gpu__time_duration) of the reference's llama-bench pp512 through the plugin: where a prefill ubatch's time goes
set -u
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out
G=/tmp/8b-q4_k_m.gguf
python scripts/make_gguf.py --model 8b --quant q4_k_m -o $G || exit 1
LB=oracle/_ref/avx2/llama-bench
GGML_BACKEND_PATH=$PWD/super-llama.cpp_b200/libggml-b200.so timeout 600 $LB -m $G -p 512 -n 0 -r 2 -t $(nproc) -fa 1 -ngl 99 2>&1 | grep pp512
GGML_BACKEND_PATH=$PWD/super-llama.cpp_b200/libggml-b200.so timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/pp_launches.csv $LB -m $G -p 512 -n 0 -r 1 -t $(nproc) -fa 1 -ngl 99 > gpurun_out/pp_launches.log 2>&1; echo "ncu rc=$?"
python - <<'PY'
import csv, collections
rows = [r for r in csv.reader(open("gpurun_out/pp_launches.csv")) if len(r) > 10 and r[0].isdigit()]
hdr = None
for r in csv.reader(open("gpurun_out/pp_launches.csv")):
    if r and r[0] == "ID": hdr = r; break
kn, mv = hdr.index("Kernel Name"), hdr.index("Metric Value")
names = [r[kn] for r in rows]; vals = [float(r[mv].replace(",", "")) for r in rows]
print(len(rows), "launches")
# the last complete ubatch: from the last k_get_rows / first kernel pattern; simply take the second half
n = len(rows); half = rows and n // 2
agg = collections.OrderedDict()
for nm, v in list(zip(names, vals))[half:]:
    k = nm.split("(")[0][:60]
    a = agg.setdefault(k, [0, 0.0]); a[0] += 1; a[1] += v
tot = sum(a[1] for a in agg.values())
for k, a in sorted(agg.items(), key=lambda x: -x[1][1]): print(f"{a[1]/1e3:9.1f} us {100*a[1]/tot:5.1f}%  x{a[0]:4d}  {k}")
print(f"total {tot/1e3:.1f} us over {n - half} launches (second half of the capture)")
PY
