# full GPU suite, default bench, launch list + full ncu capture of the current row-split kernel, medium sweep
set -x
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > gpurun_out/r16_pytest.log
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/r16_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r16_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/r16_ncu_list.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:spmm_rowsplit -s 3 -c 1 -o gpurun_out/r16_rowsplit_cfg2 python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/r16_ncu_full.log 2>&1
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu --workload cfg5_large > gpurun_out/r16_cfg5_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:spmm_rowsplit -s 3 -c 1 -o gpurun_out/r16_rowsplit_cfg5 python bench.py --steps 3 --warmup 3 --no-cpu --workload cfg5_large > gpurun_out/r16_ncu_full5.log 2>&1
timeout 2400 python scripts/sweep_medium.py > gpurun_out/r16_sweep.md 2> gpurun_out/r16_sweep.err
cat gpurun_out/r16_pytest.log; ls -la gpurun_out | tail -12
