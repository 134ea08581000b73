for wl in cfg2 cfg5_large cfg3b; do for k in 1 2; do for g in 0 2 4 8 16 32; do
echo -n "$wl k=$k G=$g: "; SPMM_B200_SMALLK_G=$g timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --workload $wl --k $k --variant rowsplit 2>&1 | grep -o '"ms_per_step": [0-9.]*, "higher'
done; done; done
