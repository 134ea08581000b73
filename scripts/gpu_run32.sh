set -x
for k in 32 64 128; do
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --workload cfg2 --k $k --variant rowsplit > gpurun_out/r32_cfg2_k$k.log 2>&1
SPMM_B200_PREFETCH_A=1 timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --workload cfg2 --k $k --variant rowsplit > gpurun_out/r32_cfg2_k${k}_pf.log 2>&1
done
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --workload cfg5_large --k 32 > gpurun_out/r32_cfg5_k32.log 2>&1
SPMM_B200_PREFETCH_A=1 timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --workload cfg5_large --k 32 > gpurun_out/r32_cfg5_k32_pf.log 2>&1
grep -H -o '"ms_per_step": [0-9.]*, "higher\|checksum_c_rank0": [0-9.]*' gpurun_out/r32_*.log
