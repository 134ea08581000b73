#!/bin/bash
# Round 2, GPU call 27: CTAs per SM of the ticketed kernel on the 1/8 shard of cfg2 (k = 32)
mkdir -p gpurun_out
timeout 300 python scripts/knob_probe.py --workload cfg2 --shard 0/8 --set TICKET_CTAS=4 --set TICKET_CTAS=3 --set TICKET_CTAS=2 --set TICKET_CTAS=4,TICKET=8 --set TICKET_CTAS=3,TICKET=8 --set TICKET_CTAS=4,TICKET=4 > gpurun_out/r2h_ctas_shard8.log 2>&1; cat gpurun_out/r2h_ctas_shard8.log
