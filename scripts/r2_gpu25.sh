#!/bin/bash
# Round 2, GPU call 25: fp32 window upload
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_f32.py -m gpu -q > gpurun_out/r2g_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r2g_pytest.log
