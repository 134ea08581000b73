#!/bin/bash
# Round 2, GPU call 1: the whole GPU suite, the default bench, and the probes of the new row-split kernels / shards / e2e slices.
mkdir -p gpurun_out
O=gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > $O/r2a_gpu.txt 2>&1
lscpu | grep -E "Model name|^CPU\(s\)|Thread|Core|Socket" > $O/r2a_cpu.txt 2>&1; free -g >> $O/r2a_cpu.txt
timeout 900 python -m pytest tests -m gpu -x -q > $O/r2a_pytest.log 2>&1; echo "pytest rc=$?" >> $O/r2a_pytest.log; tail -5 $O/r2a_pytest.log
timeout 120 python __graft_entry__.py smoke > $O/r2a_smoke.log 2>&1; tail -2 $O/r2a_smoke.log
timeout 400 python bench.py > $O/r2a_bench_default.log 2>&1; tail -c 1500 $O/r2a_bench_default.log
# row-split scheduling / staging on cfg2 (k = 32, 64, 128)
timeout 300 python scripts/knob_probe.py --workload cfg2 --set TICKET=0 --set TICKET=8 --set TICKET=4 --set TICKET=4,STAGE=1 --set TICKET=8,STAGE=2 --set TICKET=4,STAGE=1,TICKET_CTAS=3 --set TICKET=4,STAGE=1,CONST_LD=0 > $O/r2a_stage_cfg2_k32.log 2>&1; cat $O/r2a_stage_cfg2_k32.log
timeout 300 python scripts/knob_probe.py --workload cfg2 --k 64 --set TICKET=0 --set TICKET=8 --set TICKET=8,STAGE=1 > $O/r2a_stage_cfg2_k64.log 2>&1; cat $O/r2a_stage_cfg2_k64.log
timeout 300 python scripts/knob_probe.py --workload cfg2 --k 128 --set TICKET=0 --set TICKET=8 --set TICKET=8,STAGE=1 > $O/r2a_stage_cfg2_k128.log 2>&1; cat $O/r2a_stage_cfg2_k128.log
timeout 300 python scripts/knob_probe.py --workload cfg2_bw1 --set TICKET=8,STAGE=0 --set TICKET=4,STAGE=1 > $O/r2a_stage_cfg2bw1.log 2>&1; cat $O/r2a_stage_cfg2bw1.log
# the 1/8 and 1/2 shards of cfg2 (what one rank of the 8-GPU run executes)
timeout 300 python scripts/knob_probe.py --workload cfg2 --shard 0/8 --set TICKET=0 --set TICKET=8 --set TICKET=4 --set TICKET=4,STAGE=1 --set TICKET=0,CTAS_PER_SM=8 > $O/r2a_shard8.log 2>&1; cat $O/r2a_shard8.log
timeout 300 python scripts/knob_probe.py --workload cfg2 --shard 0/2 --set TICKET=0 --set TICKET=8 --set TICKET=4,STAGE=1 > $O/r2a_shard2.log 2>&1; cat $O/r2a_shard2.log
# host-buffer call: slice widths
timeout 300 python scripts/e2e_probe.py --workload cfg2 --set PIPELINE_COLS=16 --set PIPELINE_COLS=8 --set PIPELINE_COLS=8,DIRECT_STORE=0 --set PIPELINE_COLS=16,TICKET=4,STAGE=1 > $O/r2a_e2e_cfg2.log 2>&1; cat $O/r2a_e2e_cfg2.log
ls -la $O | tail -20
