#!/bin/bash
# N4 pipeline: GPU parity tests + stage timings
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_pipeline.py tests/test_abi.py -m gpu -x -q > gpurun_out/r43_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r43_pytest.log
tail -15 gpurun_out/r43_pytest.log
timeout 600 python scripts/pipeline_probe.py 8192 64 0.05 16 > gpurun_out/r43_probe_8k.log 2>&1; tail -25 gpurun_out/r43_probe_8k.log
timeout 600 python scripts/pipeline_probe.py 16384 64 0.05 16 > gpurun_out/r43_probe_16k.log 2>&1; tail -25 gpurun_out/r43_probe_16k.log
