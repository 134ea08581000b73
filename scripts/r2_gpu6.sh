#!/bin/bash
# Round 2, GPU call 6: TF32 kernel v3 (mbarrier hand-over, MMA warp), ticket kernel without the initial atomic storm
mkdir -p gpurun_out
O=gpurun_out
timeout 200 python scripts/tf32_debug.py > $O/r2f_tf32_debug.log 2>&1; echo "rc=$?" >> $O/r2f_tf32_debug.log; grep -E "mismatches|rc=" $O/r2f_tf32_debug.log
timeout 300 python -m pytest tests/test_f32.py -m gpu -q > $O/r2f_pytest_f32.log 2>&1; echo "rc=$?" >> $O/r2f_pytest_f32.log; tail -4 $O/r2f_pytest_f32.log
timeout 300 python scripts/tf32_probe.py --workload cfg4 --variants rowsplit,bcsr > $O/r2f_tf32_cfg4.log 2>&1; cat $O/r2f_tf32_cfg4.log
timeout 300 python scripts/knob_probe.py --workload cfg2 --shard 0/8 --set TICKET=-1 --set TICKET=0 > $O/r2f_shard8.log 2>&1; cat $O/r2f_shard8.log
timeout 300 python scripts/knob_probe.py --workload cfg2 --set TICKET=-1 > $O/r2f_cfg2.log 2>&1; cat $O/r2f_cfg2.log
timeout 400 ncu --set full --clock-control none --import-source on -k regex:spmm_tf32_blocked -s 5 -c 1 -o $O/r2f_ncu_tf32_cfg4 -f python scripts/tf32_probe.py --workload cfg4 --ks 128 --variants bcsr > $O/r2f_ncu_tf32.log 2>&1; tail -2 $O/r2f_ncu_tf32.log
timeout 900 python -m pytest tests -m gpu -q -k "not full_size" > $O/r2f_pytest.log 2>&1; echo "pytest rc=$?" >> $O/r2f_pytest.log; tail -4 $O/r2f_pytest.log
