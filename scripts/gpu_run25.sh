set -x
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/r25_pytest.log; cat gpurun_out/r25_pytest.log
timeout 2700 python scripts/sweep_medium.py > gpurun_out/r25_sweep.md 2> gpurun_out/r25_sweep.err
cp gpurun_out/sweep_medium.json gpurun_out/r25_sweep.json
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --workload cfg5_large > gpurun_out/r25_cfg5_large.log 2>&1
grep -c "/w" gpurun_out/r25_sweep.md
