#!/bin/bash
# N = 2 check of the default bench after the direct-store / ticket changes
mkdir -p gpurun_out
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 10 --warmup 5 > gpurun_out/r69_n2_default.log 2>&1
echo "rc=$?"
tail -c 2500 gpurun_out/r69_n2_default.log | grep -o '"value": [0-9.]*\|"ms_per_step": [0-9.]*\|"e2e": {[^}]*}\|"n_gpus": [0-9]*' | head -6
grep -i "error\|traceback\|mismatch" gpurun_out/r69_n2_default.log | head -5
