#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_pipeline.py -m gpu -q > gpurun_out/r49_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r49_pytest.log
tail -4 gpurun_out/r49_pytest.log
for n in 32 64 128; do
  timeout 300 python scripts/pipeline_probe.py 8192 $n 0.05 16 > gpurun_out/r49_probe_8k_n$n.log 2>&1; grep "^#\|sddmm attention mode |\|final\|all kernels\|end to end\|CPU" gpurun_out/r49_probe_8k_n$n.log
done
SPMM_B200_SDDMM_TUNE=23 timeout 300 python scripts/pipeline_probe.py 8192 128 0.05 16 2>&1 | grep "sddmm attention mode |"
timeout 300 python scripts/pipeline_probe.py 16384 64 0.05 16 > gpurun_out/r49_probe_16k_n64.log 2>&1; grep "^#\|sddmm attention mode |\|final\|all kernels\|end to end\|CPU" gpurun_out/r49_probe_16k_n64.log
