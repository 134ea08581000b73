set -x
nvidia-smi -L | head -2; lscpu | grep -E "Model name|^CPU\(s\)|Thread|Core|Socket" 
SPMM_B200_TEST_VARIANTS=rowsplit,auto timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r1_pytest.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/r1_smoke.log 2>&1
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r1_bench.log 2>&1
timeout 600 python bench.py --steps 20 --warmup 5 --workload cfg2_bw1 --no-cpu > gpurun_out/r1_bench_bw1.log 2>&1
timeout 600 python bench.py --steps 20 --warmup 5 --workload cfg3a --no-cpu > gpurun_out/r1_bench_3a.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/r1_ncu.log 2>&1
tail -3 gpurun_out/r1_*.log
