#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_pipeline.py -m gpu -q > gpurun_out/r50_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r50_pytest.log
tail -4 gpurun_out/r50_pytest.log
for f in 0 1; do
  echo "== SPMM_B200_PIPELINE_FORK=$f"
  SPMM_B200_PIPELINE_FORK=$f timeout 300 python scripts/pipeline_probe.py 8192 64 0.05 16 2>&1 | grep "all kernels\|end to end"
  SPMM_B200_PIPELINE_FORK=$f timeout 300 python scripts/pipeline_probe.py 16384 64 0.05 16 2>&1 | grep "all kernels\|end to end"
done > gpurun_out/r50_fork.log 2>&1
cat gpurun_out/r50_fork.log
