#!/bin/bash
mkdir -p gpurun_out
timeout 70 python -m pytest tests/test_pipeline_f32.py tests -m gpu -x -q > gpurun_out/r73_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r73_pytest.log
tail -12 gpurun_out/r73_pytest.log
