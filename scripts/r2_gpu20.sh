#!/bin/bash
# Round 2, GPU call 20 (N GPUs): routed downloads with one process per GPU (bench.py under torchrun), single-process routed test
N=${1:-2}
mkdir -p gpurun_out
O=gpurun_out
timeout 600 python -m pytest tests/test_rowband.py -m gpu -q -x -k "routed or routes or own_windows" > $O/r2p${N}_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 $O/r2p${N}_pytest.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519"
show() { python - <<PY
import json
for ln in open("$1"):
    if ln.startswith("{"):
        d = json.loads(ln); e = d["e2e"]
        print("$2", "ms", round(d["ms_per_step"], 4), "GF", round(d["value"]), "e2e_ms", round(e["ms_per_step"], 3), "parity", d.get("parity", {}).get("max_rel_err"), "|", e.get("pipeline"), "|", e.get("download_routes", {}).get("route"), e.get("download_routes", {}).get("probe_gbs_all_lower_upper_even_odd"), "| h2d", e["h2d_bytes_per_step"], "d2h", e["d2h_bytes_per_step"])
PY
}
if [ "$N" = "2" ]; then
  SPMM_B200_BENCH_FORCE_ROUTE=1,1 timeout 600 $TR bench.py --gpus $N --steps 10 --warmup 3 > $O/r2p${N}_bench_forced.log 2>&1; tail -2 $O/r2p${N}_bench_forced.log | cut -c1-200; show $O/r2p${N}_bench_forced.log "N=$N forced 1,1"
  timeout 600 $TR bench.py --gpus $N --steps 10 --warmup 3 > $O/r2p${N}_bench_cfg2.log 2>&1; show $O/r2p${N}_bench_cfg2.log "N=$N default"
else
  for wl in cfg2 cfg4; do
    timeout 600 $TR bench.py --gpus $N --workload $wl --steps 20 --warmup 5 > $O/r2p${N}_bench_$wl.log 2>&1; tail -2 $O/r2p${N}_bench_$wl.log | cut -c1-200; show $O/r2p${N}_bench_$wl.log "N=$N $wl"
  done
  SPMM_B200_BENCH_NO_ROUTE=1 timeout 600 $TR bench.py --gpus $N --steps 20 --warmup 5 > $O/r2p${N}_bench_cfg2_noroute.log 2>&1; show $O/r2p${N}_bench_cfg2_noroute.log "N=$N cfg2 no routing"
fi
