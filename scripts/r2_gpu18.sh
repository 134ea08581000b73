#!/bin/bash
# Round 2, GPU call 18 (N GPUs): download routing of spmm_b200_run_multi -- tests, in-process e2e probe, the reference harness
N=${1:-2}
mkdir -p gpurun_out
O=gpurun_out
timeout 600 python -m pytest tests/test_rowband.py tests/test_gpu_round2.py -m gpu -q -x -k "multi" > $O/r2x${N}_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 $O/r2x${N}_pytest.log
if [ "$N" = "8" ]; then SETS="--set MULTI_ROUTE=-1 --set MULTI_ROUTE=0 --set MULTI_ROUTE=240 --set MULTI_ROUTE=15 --set MULTI_ROUTE=170 --set MULTI_ROUTE=-1"; else SETS="--set MULTI_ROUTE=-1 --set MULTI_ROUTE=0 --set MULTI_ROUTE=1 --set MULTI_ROUTE=2"; fi
timeout 300 python scripts/e2e_probe.py --workload cfg2 --multi $N $SETS > $O/r2x${N}_e2e_inproc_cfg2.log 2>&1; cat $O/r2x${N}_e2e_inproc_cfg2.log
timeout 300 python scripts/e2e_probe.py --workload cfg4 --multi $N --set MULTI_ROUTE=-1 --set MULTI_ROUTE=0 > $O/r2x${N}_e2e_inproc_cfg4.log 2>&1; cat $O/r2x${N}_e2e_inproc_cfg4.log
env USE_PROCESSES=0 USE_ARTIFICIAL_MATRICES=1 USE_DLCM_MATRICES=0 NUM_COLS=32 COOLDOWN=0 RAPL_REGISTERS=0 OMP_NUM_THREADS=16 SPMM_B200_NGPUS=$N \
  timeout 600 oracle/_ref/spmm_b200_d.exe 1000000 1000000 50 16.6667 normal random 0.05 0 0.05 0.05 14 > $O/r2x${N}_harness_stdout.log 2> $O/r2x${N}_harness_csv.log; echo "harness rc=$?"; tail -1 $O/r2x${N}_harness_csv.log | cut -c1-400
