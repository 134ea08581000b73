set -x
for wl in cfg2 cfg5_large; do
for gvu in 0 222 414; do
SPMM_B200_GVU=$gvu timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --workload $wl --k 4 --variant rowsplit > gpurun_out/r36_${wl}_k4_$gvu.log 2>&1
done
for gvu in 0 424 242; do
SPMM_B200_GVU=$gvu timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --workload $wl --k 8 --variant rowsplit > gpurun_out/r36_${wl}_k8_$gvu.log 2>&1
done
done
grep -H -o '"ms_per_step": [0-9.]*, "higher\|checksum_c_rank0": [0-9.]*' gpurun_out/r36_*.log
