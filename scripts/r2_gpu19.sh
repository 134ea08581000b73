#!/bin/bash
# Round 2, GPU call 19: lane-group shape 16 x 2 (2-row tickets) on the 1/8 shard of cfg2 at k = 32
mkdir -p gpurun_out
O=gpurun_out
timeout 300 python scripts/knob_probe.py --workload cfg2 --shard 0/8 --set TICKET=4 --set SHAPE32=1,TICKET=4 --set SHAPE32=1,TICKET=2 --set SHAPE32=1,TICKET=6 --set SHAPE32=0,TICKET=4 > $O/r2q_shape_shard8.log 2>&1; cat $O/r2q_shape_shard8.log
timeout 300 python scripts/knob_probe.py --workload cfg2 --set TICKET=4 --set SHAPE32=1,TICKET=4 --set SHAPE32=1,TICKET=2 --set SHAPE32=0,TICKET=4 > $O/r2q_shape_whole.log 2>&1; cat $O/r2q_shape_whole.log
