#!/bin/bash
mkdir -p gpurun_out
timeout 600 python scripts/create_probe.py > gpurun_out/r55_create.log 2>&1; cat gpurun_out/r55_create.log
