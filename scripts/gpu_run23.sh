set -x
for wl in cfg2 cfg2_bw1 cfg5_large; do
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --workload $wl --variant rowsplit > gpurun_out/r23_$wl.log 2>&1
done
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --workload cfg2 --k 64 > gpurun_out/r23_cfg2_k64.log 2>&1
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --workload cfg2 --k 128 --variant rowsplit > gpurun_out/r23_cfg2_k128.log 2>&1
grep -h -o '"ms_per_step": [0-9.]*, "higher' gpurun_out/r23_*.log
