#!/bin/bash
# Round 2, multi-GPU call: N = $1 GPUs of one box.  In-process multi-GPU entry points (tests + the reference harness with
# SPMM_B200_NGPUS), bench.py under torchrun (value, e2e, rank-0 timeline), the in-process e2e probe, the PCIe probe.
N=${1:-2}
mkdir -p gpurun_out
O=gpurun_out
nvidia-smi topo -m > $O/r2m${N}_topo.txt 2>&1
timeout 600 python -m pytest tests/test_gpu_round2.py tests/test_rowband.py -m gpu -q -k "multi" > $O/r2m${N}_pytest_multi.log 2>&1; echo "rc=$?" >> $O/r2m${N}_pytest_multi.log; tail -5 $O/r2m${N}_pytest_multi.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
for wl in cfg2 cfg4 cfg3a; do
  timeout 600 $TR bench.py --gpus $N --workload $wl --steps 20 --warmup 5 > $O/r2m${N}_bench_$wl.log 2>&1
  python - <<PY
import json
for ln in open("$O/r2m${N}_bench_$wl.log"):
    if ln.startswith("{"):
        d = json.loads(ln); r = d["roofline"]
        e = d["e2e"]
        print("N=$N $wl", "ms", round(d["ms_per_step"], 4), "GF", round(d["value"]), "frac", round(r["frac"], 3), "e2e_ms", round(e["ms_per_step"], 3), d["variant"], "parity", d.get("parity", {}).get("max_rel_err"),
              "|", e.get("pipeline"), "|", e.get("replication"), "| h2d", e["h2d_bytes_per_step"], "d2h", e["d2h_bytes_per_step"])
        for t in d["e2e"].get("timeline_ms_last_step_rank0", []): print("    ", {k: round(v, 3) for k, v in t.items()})
PY
done
timeout 300 python scripts/e2e_probe.py --workload cfg2 --multi $N --set UPLOAD_WINDOW=1 --set UPLOAD_WINDOW=0 > $O/r2m${N}_e2e_inproc_cfg2.log 2>&1; cat $O/r2m${N}_e2e_inproc_cfg2.log
timeout 300 python scripts/e2e_probe.py --workload cfg2_bw1 --multi $N > $O/r2m${N}_e2e_inproc_bw1.log 2>&1; cat $O/r2m${N}_e2e_inproc_bw1.log
timeout 300 $TR scripts/pcie_multi_probe.py > $O/r2m${N}_pcie.log 2>&1; grep "^{" $O/r2m${N}_pcie.log
env USE_PROCESSES=0 USE_ARTIFICIAL_MATRICES=1 USE_DLCM_MATRICES=0 NUM_COLS=32 COOLDOWN=0 RAPL_REGISTERS=0 OMP_NUM_THREADS=16 SPMM_B200_NGPUS=$N \
  timeout 600 oracle/_ref/spmm_b200_d.exe 1000000 1000000 50 16.6667 normal random 0.05 0 0.05 0.05 14 > $O/r2m${N}_harness_stdout.log 2> $O/r2m${N}_harness_csv.log; echo "harness rc=$?"; tail -1 $O/r2m${N}_harness_csv.log | cut -c1-400
