#!/bin/bash
mkdir -p gpurun_out
run() { python bench.py --steps 30 --warmup 5 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('$1', 'ms_per_step', round(d['ms_per_step'],4))"; }
{
run "default"
for t in 16 64 256 1024; do SPMM_B200_TICKET=$t run "ticket=$t ctas=4"; done
SPMM_B200_TICKET=64 SPMM_B200_TICKET_CTAS=8 run "ticket=64 ctas=8"
run "default again"
} > gpurun_out/r65_ticket.log 2>&1
cat gpurun_out/r65_ticket.log
