#!/bin/bash
mkdir -p gpurun_out
{
for t in 0 23 42; do echo "== n=32 SPMM_B200_FUSED_TUNE=$t"; SPMM_B200_FUSED_TUNE=$t timeout 300 python scripts/pipeline_probe.py 8192 32 0.05 16 2>&1 | grep "row softmax"; done
for t in 0 42; do echo "== n=64 SPMM_B200_FUSED_TUNE=$t"; SPMM_B200_FUSED_TUNE=$t timeout 300 python scripts/pipeline_probe.py 8192 64 0.05 16 2>&1 | grep "row softmax"; done
for t in 0 11 12; do echo "== n=128 SPMM_B200_FUSED_TUNE=$t"; SPMM_B200_FUSED_TUNE=$t timeout 300 python scripts/pipeline_probe.py 8192 128 0.05 16 2>&1 | grep "row softmax"; done
} > gpurun_out/r52_fused.log 2>&1
cat gpurun_out/r52_fused.log
