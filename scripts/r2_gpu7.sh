#!/bin/bash
# Round 2, GPU call 7: TF32 kernel with split producer roles (2 and 3 stages), self-resetting ticket counters, graph tests
mkdir -p gpurun_out
O=gpurun_out
timeout 300 python -m pytest tests/test_f32.py tests/test_gpu_round2.py -m gpu -q -k "not full_size" > $O/r2g_pytest.log 2>&1; echo "rc=$?" >> $O/r2g_pytest.log; tail -6 $O/r2g_pytest.log
timeout 300 python scripts/tf32_probe.py --workload cfg4 --variants rowsplit,bcsr > $O/r2g_tf32_cfg4.log 2>&1; cat $O/r2g_tf32_cfg4.log
timeout 300 python scripts/tf32_probe.py --workload cfg4 --variants bcsr --set TF32_STAGES=3 > $O/r2g_tf32_cfg4_s3.log 2>&1; cat $O/r2g_tf32_cfg4_s3.log
timeout 300 python scripts/knob_probe.py --workload cfg2 --shard 0/8 --set TICKET=-1 --set TICKET=0 > $O/r2g_shard8.log 2>&1; cat $O/r2g_shard8.log
timeout 300 python scripts/knob_probe.py --workload cfg2 --set TICKET=-1 > $O/r2g_cfg2.log 2>&1; cat $O/r2g_cfg2.log
timeout 400 ncu --set full --clock-control none --import-source on -k regex:spmm_tf32_blocked -s 5 -c 1 -o $O/r2g_ncu_tf32_cfg4 -f python scripts/tf32_probe.py --workload cfg4 --ks 128 --variants bcsr > $O/r2g_ncu_tf32.log 2>&1; tail -1 $O/r2g_ncu_tf32.log
