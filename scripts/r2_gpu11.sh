#!/bin/bash
# Round 2, GPU call 11: k-slice measurement extended to k = 32 / 64 MB windows
mkdir -p gpurun_out
O=gpurun_out
for wl in cfg2_bw0.3 cfg2_bw1 cfg2; do
  timeout 300 python bench.py --workload $wl --steps 20 --warmup 5 --no-cpu > $O/r2k_bench_$wl.log 2>&1
  python - <<PY
import json
for ln in open("$O/r2k_bench_$wl.log"):
    if ln.startswith("{"):
        d = json.loads(ln); r = d["roofline"]
        print("$wl", "ms", round(d["ms_per_step"], 4), "GF", round(d["value"]), "frac", round(r["frac"], 4), "e2e_ms", round(d["e2e"]["ms_per_step"], 2), d["variant"], "kslice", d["kslice"], "parity", d["parity"]["max_rel_err"])
PY
done
timeout 300 python scripts/knob_probe.py --workload cfg2_bw0.3 --set KSLICE=-1 --set KSLICE=16 --set KSLICE=8 > $O/r2k_kslice_bw03.log 2>&1; cat $O/r2k_kslice_bw03.log
timeout 300 python scripts/knob_probe.py --workload cfg2_bw1 --set KSLICE=-1 --set KSLICE=16 --set KSLICE=8 > $O/r2k_kslice_bw1.log 2>&1; cat $O/r2k_kslice_bw1.log
timeout 300 python scripts/knob_probe.py --workload cfg2_bw0.3 --k 64 --set KSLICE=-1 --set KSLICE=32 --set KSLICE=16 > $O/r2k_kslice_bw03_k64.log 2>&1; cat $O/r2k_kslice_bw03_k64.log
