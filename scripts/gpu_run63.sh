#!/bin/bash
mkdir -p gpurun_out
{
for rev in 1 0; do for tune in 0 22; do
  echo "== TICKET_REV=$rev SDDMM_TUNE=$tune"
  for args in "8192 64" "16384 64"; do
    SPMM_B200_TICKET_REV=$rev SPMM_B200_SDDMM_TUNE=$tune timeout 300 python scripts/pipeline_probe.py $args 0.05 16 2>&1 | grep "sddmm attention mode |\|fused"
  done
done; done
} > gpurun_out/r63_rev.log 2>&1
cat gpurun_out/r63_rev.log
