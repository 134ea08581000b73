#!/bin/bash
# Round 2, GPU call 14: full GPU suite after the window-upload / row-band changes, interleaved enqueue, bench lines
mkdir -p gpurun_out
O=gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > $O/r2t_pytest.log 2>&1; echo "pytest rc=$?"; tail -8 $O/r2t_pytest.log
timeout 300 python scripts/e2e_probe.py --workload cfg2 --set ROWBAND=-1 --set ROWBAND=8 > $O/r2t_e2e_cfg2.log 2>&1; cat $O/r2t_e2e_cfg2.log
timeout 300 python bench.py --steps 20 --warmup 5 > $O/r2t_bench_default.log 2>&1; tail -c 1200 $O/r2t_bench_default.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/r2t_smoke.log 2>&1; tail -3 $O/r2t_smoke.log
# the reference's unchanged harness through the plug-in TU (host buffers, its own clock)
NUM_COLS=32 USE_PROCESSES=0 USE_ARTIFICIAL_MATRICES=1 USE_DLCM_MATRICES=0 COOLDOWN=0 RAPL_REGISTERS=0 OMP_NUM_THREADS=16 timeout 300 oracle/_ref/spmm_b200_d.exe 1000000 1000000 50 16.6667 normal random 0.05 0 0.05 0.05 14 > $O/r2t_harness_cfg2.log 2>&1; tail -3 $O/r2t_harness_cfg2.log
