#!/bin/bash
# Round 2, GPU call 23 (4 GPUs): bench.py under torchrun with the route measurement active
mkdir -p gpurun_out
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29521"
timeout 600 $TR bench.py --gpus 4 --steps 20 --warmup 5 > $O/r2p4_bench_cfg2.log 2>&1
python - <<PY
import json
for ln in open("$O/r2p4_bench_cfg2.log"):
    if ln.startswith("{"):
        d = json.loads(ln); e = d["e2e"]
        print("N=4 cfg2", "ms", round(d["ms_per_step"], 4), "GF", round(d["value"]), "e2e_ms", round(e["ms_per_step"], 3), "parity", d.get("parity", {}).get("max_rel_err"), "|", e.get("pipeline"), "|", e.get("download_routes", {}).get("route"), e.get("download_routes", {}).get("probe_gbs_all_lower_upper_even_odd"))
PY
tail -3 $O/r2p4_bench_cfg2.log | cut -c1-200
