#!/bin/bash
# Round 2, GPU call 12: the row-band host pipeline (run_rowband) -- tests, then the e2e probe on cfg2 / bw 0.3 / bw 1
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests/test_rowband.py tests/test_gpu_round2.py -m gpu -x -q -k "rowband or timeline or multi_gpu_entry or harness" > $O/r2r_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 $O/r2r_pytest.log
timeout 600 python scripts/e2e_probe.py --workload cfg2 --set ROWBAND=0 --set ROWBAND=8,DIRECT_STORE=1 --set ROWBAND=16,DIRECT_STORE=1 --set ROWBAND=32,DIRECT_STORE=1 \
   --set ROWBAND=8,DIRECT_STORE=0 --set ROWBAND=16,DIRECT_STORE=0 --set ROWBAND=32,DIRECT_STORE=0 --set ROWBAND=-1,DIRECT_STORE=1 > $O/r2r_e2e_cfg2.log 2>&1; grep -v "slice" $O/r2r_e2e_cfg2.log
timeout 600 python scripts/e2e_probe.py --workload cfg2_bw0.3 --set ROWBAND=0 --set ROWBAND=8,DIRECT_STORE=1 --set ROWBAND=16,DIRECT_STORE=0 --set ROWBAND=-1,DIRECT_STORE=1 > $O/r2r_e2e_bw03.log 2>&1; grep -v "slice" $O/r2r_e2e_bw03.log
timeout 600 python scripts/e2e_probe.py --workload cfg2_bw1 --set ROWBAND=0 --set ROWBAND=8,DIRECT_STORE=1 --set ROWBAND=-1,DIRECT_STORE=1 > $O/r2r_e2e_bw1.log 2>&1; grep -v "slice" $O/r2r_e2e_bw1.log
timeout 600 python scripts/e2e_probe.py --workload cfg2 --k 128 --set ROWBAND=0 --set ROWBAND=16,DIRECT_STORE=1 --set ROWBAND=16,DIRECT_STORE=0 --set ROWBAND=-1,DIRECT_STORE=1 > $O/r2r_e2e_cfg2_k128.log 2>&1; grep -v "slice" $O/r2r_e2e_cfg2_k128.log
timeout 300 python bench.py --steps 10 --warmup 3 > $O/r2r_bench_default.log 2>&1; tail -c 1500 $O/r2r_bench_default.log
