#!/bin/bash
mkdir -p gpurun_out
{ echo "== tickets (default)"; F32_ONLY=cfg2 timeout 200 python scripts/f32_probe.py 2>&1 | grep "cfg2"; echo "== static ranges (SPMM_B200_TICKET=0)"; SPMM_B200_TICKET=0 F32_ONLY=cfg2 timeout 200 python scripts/f32_probe.py 2>&1 | grep "cfg2"; } > gpurun_out/r70_f32_ticket.log 2>&1
cat gpurun_out/r70_f32_ticket.log
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/r70_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r70_pytest.log
tail -3 gpurun_out/r70_pytest.log
timeout 200 python __graft_entry__.py smoke 2>&1 | tail -1
