# usage: bash scripts/gpu_multi.sh N  -- bench under torchrun on N GPUs of one box
N=$1
set -x
for wl in cfg2 cfg3a cfg4 cfg5_large; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 20 --warmup 5 --workload $wl > gpurun_out/m${N}_$wl.log 2>&1
done
tail -c 600 gpurun_out/m${N}_cfg2.log
