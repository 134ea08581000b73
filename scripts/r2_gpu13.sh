#!/bin/bash
# Round 2, GPU call 13: touched-window upload + row-band pipeline (equal / graded blocks): tests, e2e probe, bench
mkdir -p gpurun_out
O=gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > $O/r2s_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 $O/r2s_pytest.log
timeout 600 python scripts/e2e_probe.py --workload cfg2 --set ROWBAND=-1 --set ROWBAND=-2 --set ROWBAND=8 --set ROWBAND=4 --set ROWBAND=0 --set ROWBAND=0,UPLOAD_WINDOW=0 --set ROWBAND=8,UPLOAD_WINDOW=0 > $O/r2s_e2e_cfg2.log 2>&1; grep -v "slice" $O/r2s_e2e_cfg2.log
timeout 600 python scripts/e2e_probe.py --workload cfg2_bw0.3 --set ROWBAND=-1 --set ROWBAND=-2 --set ROWBAND=8 --set ROWBAND=0 > $O/r2s_e2e_bw03.log 2>&1; grep -v "slice" $O/r2s_e2e_bw03.log
timeout 600 python scripts/e2e_probe.py --workload cfg2_bw1 --set ROWBAND=-1 --set ROWBAND=0 > $O/r2s_e2e_bw1.log 2>&1; grep -v "slice" $O/r2s_e2e_bw1.log
timeout 600 python scripts/e2e_probe.py --workload cfg2 --k 128 --set ROWBAND=-1 --set ROWBAND=-2 --set ROWBAND=0 > $O/r2s_e2e_cfg2_k128.log 2>&1; grep -v "slice" $O/r2s_e2e_cfg2_k128.log
timeout 600 python scripts/e2e_probe.py --workload cfg5_large --set ROWBAND=-1 --set ROWBAND=0 > $O/r2s_e2e_cfg5.log 2>&1; grep -v "slice" $O/r2s_e2e_cfg5.log
timeout 300 python bench.py --steps 20 --warmup 5 > $O/r2s_bench_default.log 2>&1; tail -c 2500 $O/r2s_bench_default.log
for wl in cfg3a cfg4; do timeout 300 python bench.py --workload $wl --steps 10 --warmup 3 --no-cpu > $O/r2s_bench_$wl.log 2>&1; python - <<PY
import json
for ln in open("$O/r2s_bench_$wl.log"):
    if ln.startswith("{"):
        d = json.loads(ln); e = d["e2e"]
        print("$wl", "ms", round(d["ms_per_step"], 4), "e2e_ms", round(e["ms_per_step"], 2), e["pipeline"], "h2d", e["h2d_bytes_per_step"], "full", e.get("full_b_upload"), "parity", d["parity"]["max_rel_err"])
PY
done
