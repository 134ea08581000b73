#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_pipeline.py -m gpu -q -x > gpurun_out/r51_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r51_pytest.log
tail -12 gpurun_out/r51_pytest.log
for t in 0 21 22 24 43; do
  echo "== SPMM_B200_FUSED_TUNE=$t"
  SPMM_B200_FUSED_TUNE=$t timeout 300 python scripts/pipeline_probe.py 8192 64 0.05 16 2>&1 | grep "row softmax"
  SPMM_B200_FUSED_TUNE=$t timeout 300 python scripts/pipeline_probe.py 16384 64 0.05 16 2>&1 | grep "row softmax"
done > gpurun_out/r51_fused.log 2>&1
cat gpurun_out/r51_fused.log
for n in 32 128; do timeout 300 python scripts/pipeline_probe.py 8192 $n 0.05 16 2>&1 | grep "^#\|row softmax"; done >> gpurun_out/r51_fused.log 2>&1; tail -6 gpurun_out/r51_fused.log
