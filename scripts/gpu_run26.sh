set -x
timeout 1500 python -m pytest tests -m gpu -x -q -k "bcsr or generated or golden or all_k or edge" 2>&1 | tail -8 > gpurun_out/r26_pytest.log; cat gpurun_out/r26_pytest.log
for k in 128 64 32; do
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --workload cfg4 --k $k > gpurun_out/r26_cfg4_k$k.log 2>&1
SPMM_B200_BCSR_BAND=0 timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --workload cfg4 --k $k > gpurun_out/r26_cfg4_k${k}_noband.log 2>&1
done
grep -h -o '"ms_per_step": [0-9.]*, "higher' gpurun_out/r26_cfg4*.log
