set -x
for wl in cfg4_bw1e-4 cfg4_bw5e-5; do
SPMM_B200_BCSR_BAND_MAX_KB=226 timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --workload $wl > gpurun_out/r27_${wl}_band.log 2>&1
SPMM_B200_BCSR_BAND=0 timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --workload $wl > gpurun_out/r27_${wl}_noband.log 2>&1
done
grep -h -o '"ms_per_step": [0-9.]*, "higher\|"bcsr_fill": [0-9.a-z]*' gpurun_out/r27_*.log
