set -x
SPMM_B200_TEST_VARIANTS=bcsr timeout 600 python -m pytest tests -m gpu -x -q -k "not full_size and not harness" 2>&1 | tail -25 > gpurun_out/r6_pytest.log
for brw in 1 2 8; do
SPMM_B200_BCSR_BRW=$brw timeout 300 python bench.py --steps 20 --warmup 5 --workload cfg4 --variant bcsr --no-cpu > gpurun_out/r6_cfg4_tma_brw$brw.log 2>&1
SPMM_B200_BCSR_BRW=$brw timeout 300 python bench.py --steps 20 --warmup 5 --workload cfg4 --k 32 --variant bcsr --no-cpu > gpurun_out/r6_cfg4k32_tma_brw$brw.log 2>&1
done
timeout 300 python bench.py --steps 20 --warmup 5 --workload cfg4 --variant bcsr --no-cpu > gpurun_out/r6_cfg4_tma_auto.log 2>&1
timeout 300 python bench.py --steps 20 --warmup 5 --workload cfg4 --k 64 --variant bcsr --no-cpu > gpurun_out/r6_cfg4k64_tma_auto.log 2>&1
timeout 300 python bench.py --steps 20 --warmup 5 --workload cfg2 --variant bcsr --no-cpu > gpurun_out/r6_cfg2_tma_auto.log 2>&1
