set -x
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > gpurun_out/r3_pytest.log
for u in 2 4 8; do for c in 8 16 32 64; do
SPMM_B200_UNROLL=$u SPMM_B200_CTAS_PER_SM=$c timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu > gpurun_out/r3_cfg2_u${u}_c${c}.log 2>&1
done; done
timeout 600 python bench.py --steps 20 --warmup 5 --workload cfg4 --no-cpu > gpurun_out/r3_cfg4.log 2>&1
timeout 600 python bench.py --steps 20 --warmup 5 --workload cfg2_bw1 --no-cpu > gpurun_out/r3_cfg2_bw1.log 2>&1
timeout 600 python bench.py --steps 20 --warmup 5 --workload cfg2_bw0.3 --no-cpu > gpurun_out/r3_cfg2_bw03.log 2>&1
