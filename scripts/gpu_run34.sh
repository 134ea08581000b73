set -x
for cps in 4 8 16 32; do
SPMM_B200_CTAS_PER_SM=$cps timeout 600 python bench.py --steps 200 --warmup 20 --no-cpu --workload cfg1 > gpurun_out/r34_cfg1_cps$cps.log 2>&1
SPMM_B200_CTAS_PER_SM=$cps timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --workload cfg2 --k 1 > gpurun_out/r34_cfg2k1_cps$cps.log 2>&1
SPMM_B200_CTAS_PER_SM=$cps timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --workload cfg2 --k 8 > gpurun_out/r34_cfg2k8_cps$cps.log 2>&1
done
grep -H -o '"ms_per_step": [0-9.]*, "higher' gpurun_out/r34_*.log
