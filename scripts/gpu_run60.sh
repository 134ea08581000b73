#!/bin/bash
mkdir -p gpurun_out
{
for c in 0 2 3 4 6; do
  echo "== SPMM_B200_CTAS_PER_SM=$c (0 = default)"
  for n in 32 64; do
    if [ $c = 0 ]; then timeout 300 python scripts/pipeline_probe.py 8192 $n 0.05 16 2>&1 | grep "spmm('K')\|spmm('final')\|mode 0, all"; else
    SPMM_B200_CTAS_PER_SM=$c timeout 300 python scripts/pipeline_probe.py 8192 $n 0.05 16 2>&1 | grep "spmm('K')\|spmm('final')\|mode 0, all"; fi
  done
done
} > gpurun_out/r60_ctas.log 2>&1
cat gpurun_out/r60_ctas.log
