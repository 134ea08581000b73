#!/bin/bash
# full GPU suite + smoke + default bench after the N4 widening
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r46_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r46_pytest.log
tail -6 gpurun_out/r46_pytest.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/r46_smoke.log 2>&1; tail -3 gpurun_out/r46_smoke.log
timeout 600 python bench.py > gpurun_out/r46_bench.log 2>&1; tail -2 gpurun_out/r46_bench.log | cut -c1-600
