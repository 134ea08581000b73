set -x
timeout 600 python bench.py --steps 2 --warmup 3 --workload cfg4 --variant bcsr --no-cpu > gpurun_out/r5_plain_a.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:spmm_bcsr -s 3 -c 1 -o gpurun_out/r5_bcsr_dmma_cfg4 python bench.py --steps 2 --warmup 3 --workload cfg4 --variant bcsr --no-cpu > gpurun_out/r5_ncu_a.log 2>&1
timeout 600 python bench.py --steps 2 --warmup 3 --workload cfg4 --variant bcsr_fma --no-cpu > gpurun_out/r5_plain_b.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:spmm_bcsr -s 3 -c 1 -o gpurun_out/r5_bcsr_fma_cfg4 python bench.py --steps 2 --warmup 3 --workload cfg4 --variant bcsr_fma --no-cpu > gpurun_out/r5_ncu_b.log 2>&1
