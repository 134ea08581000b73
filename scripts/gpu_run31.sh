set -x
timeout 900 python -m pytest tests -m gpu -x -q -k "all_k or golden or edge or generated or shards" 2>&1 | tail -4
for k in 32 64 128 16; do
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --workload cfg2 --k $k --variant rowsplit > gpurun_out/r31_cfg2_k$k.log 2>&1
SPMM_B200_CONST_LD=0 timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --workload cfg2 --k $k --variant rowsplit > gpurun_out/r31_cfg2_k${k}_old.log 2>&1
done
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --workload cfg5_large --k 32 > gpurun_out/r31_cfg5_k32.log 2>&1
SPMM_B200_CONST_LD=0 timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --workload cfg5_large --k 32 > gpurun_out/r31_cfg5_k32_old.log 2>&1
grep -H -o '"ms_per_step": [0-9.]*, "higher' gpurun_out/r31_*.log
