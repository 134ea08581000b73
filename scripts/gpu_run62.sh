#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_pipeline.py -m gpu -q > gpurun_out/r62_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r62_pytest.log
tail -4 gpurun_out/r62_pytest.log
{
for args in "8192 32" "8192 64" "8192 128" "16384 64"; do
  timeout 300 python scripts/pipeline_probe.py $args 0.05 16 2>&1 | grep "^#\|sddmm attention mode |\|spmm('final')\|all kernels\|fused\|end to end"
done
} > gpurun_out/r62_ticket.log 2>&1
cat gpurun_out/r62_ticket.log
