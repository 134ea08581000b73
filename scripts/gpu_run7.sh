set -x
timeout 300 python bench.py --steps 2 --warmup 3 --workload cfg4 --variant bcsr --no-cpu > gpurun_out/r7_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:spmm_bcsr -s 3 -c 1 -o gpurun_out/r7_bcsr_tma_cfg4 python bench.py --steps 2 --warmup 3 --workload cfg4 --variant bcsr --no-cpu > gpurun_out/r7_ncu.log 2>&1
