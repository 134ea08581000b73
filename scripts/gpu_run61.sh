#!/bin/bash
mkdir -p gpurun_out
{
for args in "8192 64" "8192 128" "16384 64" "4096 64"; do
  timeout 300 python scripts/pipeline_probe.py $args 0.05 16 2>&1 | grep "^#\|spmm('K')\|spmm('final')\|all kernels\|fused"
done
} > gpurun_out/r61_wave.log 2>&1
cat gpurun_out/r61_wave.log
