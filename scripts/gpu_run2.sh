# full GPU test suite + merge-path bench + one full ncu capture of the row-split kernel
set -x
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -25 > gpurun_out/r2_pytest.log
timeout 600 python bench.py --steps 20 --warmup 5 --workload cfg3a --no-cpu > gpurun_out/r2_bench_3a.log 2>&1
timeout 600 python bench.py --steps 20 --warmup 5 --workload cfg3b --no-cpu > gpurun_out/r2_bench_3b.log 2>&1
timeout 600 python bench.py --steps 20 --warmup 5 --workload cfg3b --variant merge --no-cpu > gpurun_out/r2_bench_3b_merge.log 2>&1
timeout 600 python bench.py --steps 20 --warmup 5 --workload cfg2 --variant merge --no-cpu > gpurun_out/r2_bench_cfg2_merge.log 2>&1
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/r2_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:spmm_rowsplit -s 3 -c 1 -o gpurun_out/r2_rowsplit_cfg2 python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/r2_ncu.log 2>&1
ls -la gpurun_out
