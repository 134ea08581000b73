#!/bin/bash
# Round 2, GPU call 4: TF32 kernel after the K-major fix, revert check of the streaming hints, full GPU suite.
mkdir -p gpurun_out
O=gpurun_out
timeout 60 scripts/umma_probe.bin > $O/r2d_umma_probe.log 2>&1; echo "rc=$?" >> $O/r2d_umma_probe.log; grep -E "variant= (0|4)|N= (32|64)|rc=" $O/r2d_umma_probe.log
timeout 200 python scripts/tf32_debug.py > $O/r2d_tf32_debug.log 2>&1; echo "rc=$?" >> $O/r2d_tf32_debug.log; grep -E "mismatches|rc=" $O/r2d_tf32_debug.log
timeout 300 python scripts/tf32_probe.py --workload cfg4 > $O/r2d_tf32_cfg4.log 2>&1; cat $O/r2d_tf32_cfg4.log
timeout 300 python scripts/tf32_probe.py --workload cfg2 --ks 32 --variants rowsplit,bcsr > $O/r2d_tf32_cfg2.log 2>&1; cat $O/r2d_tf32_cfg2.log
for wl in cfg2_bw0.3 cfg2_bw1 cfg3a cfg3b; do
  timeout 300 python bench.py --workload $wl --steps 20 --warmup 5 --no-cpu --no-parity > $O/r2d_bench_$wl.log 2>&1
  python - <<PY
import json
for ln in open("$O/r2d_bench_$wl.log"):
    if ln.startswith("{"):
        d = json.loads(ln); r = d["roofline"]
        print("$wl", "ms", round(d["ms_per_step"], 4), "GF", round(d["value"]), "frac", round(r["frac"], 4), "e2e_ms", round(d["e2e"]["ms_per_step"], 2), d["variant"])
PY
done
timeout 900 python -m pytest tests -m gpu -q > $O/r2d_pytest.log 2>&1; echo "pytest rc=$?" >> $O/r2d_pytest.log; tail -6 $O/r2d_pytest.log
