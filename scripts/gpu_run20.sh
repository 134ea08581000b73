set -x
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 > gpurun_out/r20_pytest.log
for wl in cfg2 cfg3a cfg3b cfg2_bw0.3 cfg2_bw1; do
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --workload $wl > gpurun_out/r20_$wl.log 2>&1
done
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --workload cfg3b --variant rowsplit > gpurun_out/r20_cfg3b_rowsplit.log 2>&1
cat gpurun_out/r20_pytest.log
grep -h -o '"ms_per_step": [0-9.]*, "higher\|"variant": "[A-Z0-9_]*"\|"workload": "cfg[0-9a-z_.]*' gpurun_out/r20_cfg*.log
