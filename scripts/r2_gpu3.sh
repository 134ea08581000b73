#!/bin/bash
# Round 2, GPU call 3: tcgen05 descriptor probe + TF32 kernel debug, streaming-hint effect on the large-window workloads,
# ncu capture of the staged (TMA) row-split experiment for the record.
mkdir -p gpurun_out
O=gpurun_out
timeout 60 scripts/umma_probe.bin > $O/r2c_umma_probe.log 2>&1; echo "rc=$?" >> $O/r2c_umma_probe.log; cat $O/r2c_umma_probe.log
timeout 200 python scripts/tf32_debug.py > $O/r2c_tf32_debug.log 2>&1; echo "rc=$?" >> $O/r2c_tf32_debug.log; cat $O/r2c_tf32_debug.log
timeout 300 python -m pytest tests/test_f32.py -m gpu -q > $O/r2c_pytest_f32.log 2>&1; echo "rc=$?" >> $O/r2c_pytest_f32.log; tail -12 $O/r2c_pytest_f32.log
for wl in cfg2 cfg2_bw0.3 cfg2_bw1 cfg3a cfg3b cfg4 cfg5_large; do
  timeout 300 python bench.py --workload $wl --steps 20 --warmup 5 --no-cpu --no-parity > $O/r2c_bench_$wl.log 2>&1
  python - <<PY
import json
for ln in open("$O/r2c_bench_$wl.log"):
    if ln.startswith("{"):
        d = json.loads(ln); r = d["roofline"]
        print("$wl", "ms", round(d["ms_per_step"], 4), "GF", round(d["value"]), "frac", round(r["frac"], 4), "e2e_ms", round(d["e2e"]["ms_per_step"], 2), d["variant"], "kslice", d["kslice"])
PY
done
SPMM_B200_STAGE=1 SPMM_B200_TICKET=4 timeout 400 ncu --set full --clock-control none --import-source on -k regex:spmm_rowsplit_staged -s 3 -c 1 -o $O/r2c_ncu_staged_cfg2 -f python bench.py --steps 3 --warmup 3 --no-cpu --no-parity > $O/r2c_ncu_staged.log 2>&1
timeout 400 ncu --set full --clock-control none --import-source on -k regex:spmm_rowsplit_ticket -s 3 -c 1 -o $O/r2c_ncu_ticket_cfg2 -f python bench.py --steps 3 --warmup 3 --no-cpu --no-parity > $O/r2c_ncu_ticket.log 2>&1
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file $O/r2c_launches_cfg2.csv python bench.py --steps 3 --warmup 3 --no-cpu --no-parity > $O/r2c_ncu_list.log 2>&1
ls -la $O | tail -12
