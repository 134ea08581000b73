#!/bin/bash
# Round 2, GPU call 10: L2 eviction-priority hints on the large-window workloads
mkdir -p gpurun_out
O=gpurun_out
timeout 600 python scripts/knob_probe.py --workload cfg5_large --calls 10 --set KSLICE=32,L2_HINT=0 --set KSLICE=32,L2_HINT=1 --set KSLICE=16,L2_HINT=0 --set KSLICE=16,L2_HINT=1 --set KSLICE=64,L2_HINT=1 > $O/r2j_l2hint_cfg5.log 2>&1; cat $O/r2j_l2hint_cfg5.log
timeout 600 python scripts/knob_probe.py --gen "8388607 8388607 5 1.6667 normal random 0.3 0 0.05 0.05 14" --k 128 --calls 10 --set KSLICE=32,L2_HINT=0 --set KSLICE=32,L2_HINT=1 --set KSLICE=16,L2_HINT=1 > $O/r2j_l2hint_bw03.log 2>&1; cat $O/r2j_l2hint_bw03.log
timeout 600 python scripts/knob_probe.py --workload cfg5_large --k 32 --calls 10 --set TICKET=0,L2_HINT=0 > $O/r2j_cfg5_k32.log 2>&1; cat $O/r2j_cfg5_k32.log
