#!/bin/bash
# Round 2, GPU call 8: TF32 kernel (early queue refill + L2 prefetch), then the medium-dataset sweep with whole-output
# parity and the CPU column.
mkdir -p gpurun_out
O=gpurun_out
timeout 300 python -m pytest tests/test_f32.py -m gpu -q > $O/r2h_pytest_f32.log 2>&1; echo "rc=$?" >> $O/r2h_pytest_f32.log; tail -3 $O/r2h_pytest_f32.log
timeout 300 python scripts/tf32_probe.py --workload cfg4 --variants rowsplit,bcsr > $O/r2h_tf32_cfg4.log 2>&1; cat $O/r2h_tf32_cfg4.log
timeout 400 ncu --set full --clock-control none --import-source on -k regex:spmm_tf32_blocked -s 5 -c 1 -o $O/r2h_ncu_tf32_cfg4 -f python scripts/tf32_probe.py --workload cfg4 --ks 128 --variants bcsr > $O/r2h_ncu_tf32.log 2>&1; tail -1 $O/r2h_ncu_tf32.log
timeout 2400 python scripts/sweep_medium.py > $O/r2h_sweep_medium.md 2> $O/r2h_sweep_medium.err; echo "sweep rc=$?"; tail -3 $O/r2h_sweep_medium.md; tail -3 $O/r2h_sweep_medium.err
cp $O/sweep_medium.json $O/r2h_sweep_medium.json 2>/dev/null
