#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r67_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r67_pytest.log
tail -4 gpurun_out/r67_pytest.log
SWEEP_VARIANTS=rowsplit SWEEP_KS=4,8,32,128 timeout 600 python scripts/sweep_medium.py > gpurun_out/r67_sweep_ticket.md 2>&1
SPMM_B200_TICKET=0 SWEEP_VARIANTS=rowsplit SWEEP_KS=4,8,32,128 timeout 600 python scripts/sweep_medium.py > gpurun_out/r67_sweep_static.md 2>&1
python - <<'PY'
def rows(f):
    out = {}
    for l in open(f):
        p = [x.strip() for x in l.split("|")]
        if len(p) > 9 and p[1].isdigit():
            out[(p[1], p[2], p[3], p[4], p[6])] = float(p[8])
    return out
a, b = rows("gpurun_out/r67_sweep_ticket.md"), rows("gpurun_out/r67_sweep_static.md")
for k in a:
    if k in b: print(k, "static", b[k], "ticket", a[k], "ratio %.3f" % (a[k] / b[k]))
PY
