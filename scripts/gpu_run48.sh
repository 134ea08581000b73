#!/bin/bash
mkdir -p gpurun_out
for t in 0 42 43 22 23 24 82 83; do
  echo "== SPMM_B200_SDDMM_TUNE=$t"
  SPMM_B200_SDDMM_TUNE=$t timeout 300 python scripts/pipeline_probe.py 8192 64 0.05 16 2>&1 | grep "sddmm attention mode |\|mode 1, all"
  SPMM_B200_SDDMM_TUNE=$t timeout 300 python scripts/pipeline_probe.py 16384 64 0.05 16 2>&1 | grep "sddmm attention mode |"
done > gpurun_out/r48_sddmm_tune.log 2>&1
cat gpurun_out/r48_sddmm_tune.log
