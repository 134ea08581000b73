#!/bin/bash
mkdir -p gpurun_out
run() { python bench.py --steps 30 --warmup 5 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('$1', 'ms_per_step', round(d['ms_per_step'],4))"; }
{
for t in 4 8 16 32; do SPMM_B200_TICKET=$t run "ticket=$t ctas=4"; done
SPMM_B200_TICKET=8 SPMM_B200_TICKET_CTAS=8 run "ticket=8 ctas=8"
SPMM_B200_TICKET=16 SPMM_B200_TICKET_CTAS=3 run "ticket=16 ctas=3"
} > gpurun_out/r66_ticket.log 2>&1
cat gpurun_out/r66_ticket.log
