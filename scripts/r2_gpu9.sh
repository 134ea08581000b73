#!/bin/bash
# Round 2, GPU call 9: TF32 kernel v5 (chunk-major copy of A)
mkdir -p gpurun_out
O=gpurun_out
timeout 200 python scripts/tf32_debug.py > $O/r2i_tf32_debug.log 2>&1; echo "rc=$?" >> $O/r2i_tf32_debug.log; grep -E "mismatches|rc=" $O/r2i_tf32_debug.log
timeout 300 python -m pytest tests/test_f32.py -m gpu -q > $O/r2i_pytest_f32.log 2>&1; echo "rc=$?" >> $O/r2i_pytest_f32.log; tail -5 $O/r2i_pytest_f32.log
timeout 300 python scripts/tf32_probe.py --workload cfg4 --variants rowsplit,bcsr > $O/r2i_tf32_cfg4.log 2>&1; cat $O/r2i_tf32_cfg4.log
timeout 400 ncu --set full --clock-control none --import-source on -k regex:spmm_tf32_blocked -s 5 -c 1 -o $O/r2i_ncu_tf32_cfg4 -f python scripts/tf32_probe.py --workload cfg4 --ks 128 --variants bcsr > $O/r2i_ncu_tf32.log 2>&1; tail -1 $O/r2i_ncu_tf32.log
