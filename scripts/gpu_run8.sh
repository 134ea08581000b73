set -x
SPMM_B200_TEST_VARIANTS=bcsr timeout 600 python -m pytest tests -m gpu -x -q -k "not full_size and not harness" 2>&1 | tail -5 > gpurun_out/r8_pytest.log
for st in 2 3 4; do for brw in 1 2 4 8; do
SPMM_B200_BCSR_STAGES=$st SPMM_B200_BCSR_BRW=$brw timeout 300 python bench.py --steps 20 --warmup 5 --workload cfg4 --variant bcsr --no-cpu > gpurun_out/r8_cfg4_s${st}_b$brw.log 2>&1
done; done
for st in 2 4; do for brw in 2 8; do
SPMM_B200_BCSR_STAGES=$st SPMM_B200_BCSR_BRW=$brw timeout 300 python bench.py --steps 20 --warmup 5 --workload cfg4 --k 32 --variant bcsr --no-cpu > gpurun_out/r8_cfg4k32_s${st}_b$brw.log 2>&1
done; done
