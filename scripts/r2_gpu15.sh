#!/bin/bash
# Round 2, GPU call 15: poisoned-B tests over every kernel family, full suite, default bench
mkdir -p gpurun_out
O=gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > $O/r2u_pytest.log 2>&1; echo "pytest rc=$?"; tail -12 $O/r2u_pytest.log
timeout 300 python bench.py --steps 20 --warmup 5 > $O/r2u_bench_default.log 2>&1; tail -c 600 $O/r2u_bench_default.log
