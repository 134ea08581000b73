#!/bin/bash
# Round 2, GPU call 21: the allocation-free entry point -- its test alone in a fresh process, then the whole GPU suite
mkdir -p gpurun_out
O=gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > $O/r2n_pytest_alone.log 2>&1; echo "alone rc=$?"; tail -15 $O/r2n_pytest_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q > $O/r2n_pytest_all.log 2>&1; echo "all rc=$?"; tail -8 $O/r2n_pytest_all.log
