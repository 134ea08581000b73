#!/bin/bash
# Round 2, GPU call 16: row-band pipeline on the blocked variants (cfg4), tests
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests/test_rowband.py tests/test_gpu_round2.py -m gpu -q -x > $O/r2v_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 $O/r2v_pytest.log
timeout 600 python scripts/e2e_probe.py --workload cfg4 --set ROWBAND=-1 --set ROWBAND=0 --set ROWBAND=8 --set ROWBAND=16 --set ROWBAND=32 > $O/r2v_e2e_cfg4.log 2>&1; grep -v "slice" $O/r2v_e2e_cfg4.log; grep -A18 "ROWBAND=16" $O/r2v_e2e_cfg4.log | head -20
timeout 300 python bench.py --workload cfg4 --steps 10 --warmup 3 --no-cpu > $O/r2v_bench_cfg4.log 2>&1; python - <<PY
import json
for ln in open("$O/r2v_bench_cfg4.log"):
    if ln.startswith("{"):
        d = json.loads(ln); e = d["e2e"]
        print("cfg4", "ms", round(d["ms_per_step"], 4), "e2e_ms", round(e["ms_per_step"], 2), e["pipeline"], "h2d", e["h2d_bytes_per_step"], "parity", d["parity"]["max_rel_err"])
PY
