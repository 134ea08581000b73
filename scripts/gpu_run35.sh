set -x
for k in 1 2 4 8; do for v in rowsplit merge; do
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --workload cfg2 --k $k --variant $v > gpurun_out/r35_cfg2_k${k}_$v.log 2>&1
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --workload cfg5_large --k $k --variant $v > gpurun_out/r35_cfg5_k${k}_$v.log 2>&1
done; done
grep -H -o '"ms_per_step": [0-9.]*, "higher' gpurun_out/r35_*.log
