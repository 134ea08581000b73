#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/r72_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r72_pytest.log
tail -3 gpurun_out/r72_pytest.log
