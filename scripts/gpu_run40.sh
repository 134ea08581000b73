timeout 900 python -m pytest tests -m gpu -x -q -k "all_k or edge or golden or generated" 2>&1 | tail -3
for wl in cfg2 cfg5_large cfg3b cfg1; do for k in 1 2; do for r in 1 2 4; do
echo -n "$wl k=$k R=$r: "; SPMM_B200_SMALLK_R=$r timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --workload $wl --k $k --variant rowsplit 2>&1 | grep -o '"ms_per_step": [0-9.]*, "higher'
done; done; done
