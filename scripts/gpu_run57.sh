#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "direct_store or k_sliced or harness" > gpurun_out/r58_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r58_pytest.log
tail -5 gpurun_out/r58_pytest.log
