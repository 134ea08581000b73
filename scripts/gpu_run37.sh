set -x
timeout 900 python -m pytest tests -m gpu -x -q -k "all_k or edge or golden" 2>&1 | tail -3
for wl in cfg2 cfg5_large; do
for gvu in 0 212; do
SPMM_B200_GVU=$gvu timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --workload $wl --k 2 --variant rowsplit > gpurun_out/r37_${wl}_k2_$gvu.log 2>&1
done
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --workload $wl --k 4 --variant rowsplit > gpurun_out/r37_${wl}_k4.log 2>&1
done
grep -H -o '"ms_per_step": [0-9.]*, "higher' gpurun_out/r37_*.log
