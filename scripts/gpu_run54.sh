#!/bin/bash
mkdir -p gpurun_out
timeout 900 python scripts/pipeline_harness_compare.py 8192 64 > gpurun_out/r54_harness_8k.log 2>&1; cat gpurun_out/r54_harness_8k.log | cut -c1-700
timeout 900 python scripts/pipeline_harness_compare.py 16384 64 > gpurun_out/r54_harness_16k.log 2>&1; cat gpurun_out/r54_harness_16k.log | cut -c1-700
