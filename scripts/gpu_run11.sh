set -x
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r11_reference_arm.log 2>&1
timeout 600 python bench.py --steps 20 --warmup 5 --workload cfg5_large --no-cpu > gpurun_out/r11_cfg5_large.log 2>&1
timeout 600 python bench.py --steps 20 --warmup 5 --workload cfg1 > gpurun_out/r11_cfg1.log 2>&1
timeout 2400 python scripts/sweep_medium.py > gpurun_out/r11_sweep.md 2> gpurun_out/r11_sweep.err
