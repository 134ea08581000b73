set -x
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r10_pytest.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/r10_smoke.log 2>&1
