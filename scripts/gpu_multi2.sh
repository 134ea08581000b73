# usage: bash scripts/gpu_multi2.sh N  -- default bench (cfg2) + other workloads under torchrun on N GPUs
N=$1
set -x
for wl in cfg2 cfg3a cfg4 cfg5_large; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 20 --warmup 5 --workload $wl > gpurun_out/n${N}_$wl.log 2>&1
tail -c 1500 gpurun_out/n${N}_$wl.log | grep -o '"value": [0-9.]*\|"ms_per_step": [0-9.]*\|"e2e": {[^}]*}' | head -4
done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus $N --impl reference --steps 3 --warmup 1 > gpurun_out/n${N}_reference.log 2>&1
tail -c 400 gpurun_out/n${N}_reference.log
