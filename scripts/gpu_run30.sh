set -x
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 > gpurun_out/r30_pytest.log; cat gpurun_out/r30_pytest.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/r30_smoke.log 2>&1; tail -2 gpurun_out/r30_smoke.log
timeout 600 python bench.py > gpurun_out/r30_default.log 2>&1
timeout 600 python bench.py --impl reference > gpurun_out/r30_reference.log 2>&1
for wl in cfg3a cfg3b cfg2_bw0.3 cfg2_bw1; do timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --workload $wl > gpurun_out/r30_$wl.log 2>&1; done
grep -h -o '"ms_per_step": [0-9.]*, "higher' gpurun_out/r30_*.log
