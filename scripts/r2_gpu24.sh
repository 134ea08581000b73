#!/bin/bash
# Round 2, GPU call 24: fp32 softmax of the "_f" pipeline
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_pipeline_f32.py tests/test_pipeline.py -m gpu -q > gpurun_out/r2f_pytest.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/r2f_pytest.log
