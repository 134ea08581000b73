SPMM_B200_STREAM=1 timeout 900 python -m pytest tests -m gpu -x -q -k "all_k or edge or golden or generated" 2>&1 | tail -3
for wl in cfg2 cfg5_large cfg3b cfg1 cfg2_bw1; do for k in 1 2; do for st in 0 1; do
echo -n "$wl k=$k stream=$st: "; SPMM_B200_STREAM=$st timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --workload $wl --k $k --variant rowsplit 2>&1 | grep -o '"ms_per_step": [0-9.]*, "higher'
done; done; done
