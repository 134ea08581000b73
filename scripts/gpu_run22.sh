set -x
timeout 600 python -m pytest tests -m gpu -x -q -k "f_build" 2>&1 | tail -5
timeout 600 python scripts/f32_probe.py > gpurun_out/r22_f32_probe.md 2> gpurun_out/r22_err.log; cat gpurun_out/r22_f32_probe.md; tail -3 gpurun_out/r22_err.log
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/r22_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:spmm_rowsplit -s 3 -c 1 -o gpurun_out/r22_rowsplit_cfg2_g8v4 python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/r22_ncu_full.log 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r22_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/r22_ncu_list.log 2>&1
