#!/bin/bash
# ncu evidence for the default bench after the ticketed row-split kernel became the default
mkdir -p gpurun_out
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/r71_plain.log 2>&1 || { tail -5 gpurun_out/r71_plain.log; exit 1; }
timeout 400 ncu --set full --clock-control none --import-source on -k regex:spmm_rowsplit_ticket -s 3 -c 1 -o gpurun_out/r71_rowsplit_ticket_cfg2 -f python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/r71_ncu_full.log 2>&1
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/r71_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/r71_ncu_list.log 2>&1
ls -la gpurun_out/r71*
