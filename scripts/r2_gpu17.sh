#!/bin/bash
# Round 2, GPU call 17 (8 GPUs): which subsets of the 8 ranks can download at once without the aggregate collapsing
mkdir -p gpurun_out
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29517"
timeout 300 $TR scripts/pcie_multi_probe.py --mbs 32 --subsets 0,01,0123,4567,0246,1357,0145,012345,01234567,04,0426 > $O/r2w_pcie_subsets.log 2>&1; grep "^{" $O/r2w_pcie_subsets.log
