set -x
timeout 900 python -m pytest tests -m gpu -x -q -k "real_matrix or pipeline or coo" 2>&1 | tail -25 > gpurun_out/r15_pytest.log
cat gpurun_out/r15_pytest.log
