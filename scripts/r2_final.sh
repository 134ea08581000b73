#!/bin/bash
# Round 2, final validation on one GPU: whole GPU suite, smoke, both bench arms, ncu --set full + launch list of the default bench.
mkdir -p gpurun_out
O=gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > $O/r2y_pytest.log 2>&1; echo "pytest rc=$?" >> $O/r2y_pytest.log; tail -5 $O/r2y_pytest.log
timeout 120 python __graft_entry__.py smoke > $O/r2y_smoke.log 2>&1; tail -2 $O/r2y_smoke.log
timeout 600 python bench.py --impl reference > $O/r2y_bench_reference.log 2>&1; tail -c 700 $O/r2y_bench_reference.log; echo
timeout 600 python bench.py > $O/r2y_bench_default.log 2>&1; tail -c 2600 $O/r2y_bench_default.log | head -c 1500; echo
timeout 400 ncu --set full --clock-control none --import-source on -k regex:spmm_rowsplit_ticket -s 3 -c 1 -o $O/r2y_ncu_ticket_cfg2 -f python bench.py --steps 3 --warmup 3 --no-cpu --no-parity > $O/r2y_ncu_full.log 2>&1
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file $O/r2y_launches_cfg2.csv python bench.py --steps 3 --warmup 3 --no-cpu --no-parity > $O/r2y_ncu_list.log 2>&1
for wl in cfg3a cfg4 cfg5_large; do
  timeout 400 python bench.py --workload $wl --steps 20 --warmup 5 > $O/r2y_bench_$wl.log 2>&1
  python - <<PY
import json
for ln in open("$O/r2y_bench_$wl.log"):
    if ln.startswith("{"):
        d = json.loads(ln); r = d["roofline"]
        print("$wl", "ms", round(d["ms_per_step"], 4), "GF", round(d["value"]), "frac", round(r["frac"], 4), "traffic", r["traffic"], "e2e_ms", round(d["e2e"]["ms_per_step"], 2), d["variant"], "kslice", d["kslice"], "parity", d["parity"]["max_rel_err"], "cpu", round(d["cpu_baseline"]["value"], 1))
PY
done
