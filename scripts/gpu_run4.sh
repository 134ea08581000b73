set -x
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -25 > gpurun_out/r4_pytest.log
for v in rowsplit bcsr bcsr_fma; do
timeout 600 python bench.py --steps 20 --warmup 5 --workload cfg4 --variant $v --no-cpu > gpurun_out/r4_cfg4_$v.log 2>&1
timeout 600 python bench.py --steps 20 --warmup 5 --workload cfg4 --k 32 --variant $v --no-cpu > gpurun_out/r4_cfg4k32_$v.log 2>&1
timeout 600 python bench.py --steps 20 --warmup 5 --workload cfg2 --variant $v --no-cpu > gpurun_out/r4_cfg2_$v.log 2>&1
done
timeout 600 python bench.py --steps 20 --warmup 5 --workload cfg3b --no-cpu > gpurun_out/r4_cfg3b_auto.log 2>&1
