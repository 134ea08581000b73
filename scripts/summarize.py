#!/usr/bin/env python
"""Print one line per bench.py log (gpurun_out/*.log): variant, ms/step, GFLOP/s, roofline fraction."""
import json, sys
for f in sys.argv[1:]:
    try:
        d = json.loads([x for x in open(f) if x.startswith('{')][-1])
        r = d.get('roofline', {})
        print(f"{f.split('/')[-1]:40s} {d.get('variant', d.get('impl','')):20s} n={d['n_gpus']} ms={d['ms_per_step']:.4f} "
              f"GF={d['value']:.1f} frac={r.get('frac', 0):.4f} e2e={d['e2e']['value']:.1f} launches={d.get('gpu_launches')}")
    except Exception as e:
        print(f, 'ERR', e, open(f).read()[-300:])
