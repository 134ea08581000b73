# k-sliced host pipeline: parity + e2e on cfg2 / cfg4 / cfg3a
set -x
timeout 900 python -m pytest tests -m gpu -x -q -k "pipeline or full_size or harness" 2>&1 | tail -15 > gpurun_out/r13_pytest.log
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu > gpurun_out/r13_cfg2.log 2>&1
SPMM_B200_PIPELINE_COLS=0 timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu > gpurun_out/r13_cfg2_nopipe.log 2>&1
SPMM_B200_PIPELINE_COLS=8 timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu > gpurun_out/r13_cfg2_cols8.log 2>&1
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --workload cfg4 > gpurun_out/r13_cfg4.log 2>&1
SPMM_B200_PIPELINE_COLS=32 timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --workload cfg4 > gpurun_out/r13_cfg4_cols32.log 2>&1
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --workload cfg3a > gpurun_out/r13_cfg3a.log 2>&1
grep -h -o '"e2e": {[^}]*}' gpurun_out/r13_*.log
