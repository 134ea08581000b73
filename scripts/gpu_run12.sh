# re-entry verification: full GPU suite, smoke, default bench (both arms), PCIe probe for the e2e pipeline
set -x
nvidia-smi -L | head -2; lscpu | grep -E "Model name|^CPU\(s\)|Thread|Core|Socket"
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -25 > gpurun_out/r12_pytest.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/r12_smoke.log 2>&1
timeout 600 python bench.py > gpurun_out/r12_bench.log 2>&1
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r12_reference_arm.log 2>&1
nvcc -O2 -gencode arch=compute_100a,code=sm_100a -o /tmp/pcie_probe scripts/pcie_probe.cu && timeout 120 /tmp/pcie_probe > gpurun_out/r12_pcie_probe.log 2>&1
tail -4 gpurun_out/r12_*.log
