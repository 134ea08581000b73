"""Generates tests/golden/pipeline_reference.npz: outputs of the REFERENCE's own pipeline TU
(pipeline_code_bench/sddmm_taco_naive.cpp compiled into oracle/_ref/libpipeline_ref.so, its mkl_dcsrmm calls
served by the contract stand-in of oracle/ref_shim_pipeline.cpp) on small seeded inputs.  Run in the build
container (needs /root/reference):   python tests/golden/make_pipeline_golden.py
The inputs are rebuilt from the seeds by tests/pipeline_util.py; only the outputs are stored."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
sys.path.insert(0, os.path.join(HERE, ".."))
from oracle import loader as L                      # noqa: E402
from pipeline_util import pipeline_case             # noqa: E402

CASES = [  # seed, m, kdim, n, band, fill, mask values, softmax
    (1, 96, 80, 32, 4, 0.10, "ones", 0),
    (2, 200, 150, 64, 8, 0.05, "ones", 0),
    (3, 64, 64, 16, 2, 0.20, "random", 0),
    (4, 128, 100, 7, 3, 0.08, "random", 0),        # odd dense width
    (5, 150, 90, 32, 4, 0.10, "ones", 1),          # with the reference's (global) softmax
]

out = {"ncases": len(CASES)}
for i, (seed, m, kdim, n, band, fill, mv, sm) in enumerate(CASES):
    mask, W, x = pipeline_case(seed, m, kdim, n, band, fill, mask_values=mv)
    r = L.pipeline_ref(mask, W, x, n, kdim, softmax=sm)
    out[f"case{i}_params"] = np.array([seed, m, kdim, n, band, sm], dtype=np.int64)
    out[f"case{i}_fill"] = fill
    out[f"case{i}_mask_values"] = mv
    out[f"case{i}_mask_nnz"] = len(mask[1])
    for k, v in r.items():
        out[f"case{i}_{k}"] = v
np.savez_compressed(os.path.join(HERE, "pipeline_reference.npz"), **out)
print("wrote", len(CASES), "cases")

# the "_f" build of the same TU (ValueType = float): oracle/_ref/libpipeline_ref_f.so
out32 = {"ncases": 0}
for i, (seed, m, kdim, n, band, fill, mv, sm) in enumerate(c for c in CASES if c[-1] == 0):
    mask, W, x = pipeline_case(seed, m, kdim, n, band, fill, mask_values=mv)
    mask = (mask[0], mask[1], mask[2].astype(np.float32), mask[3])
    W = {t: (W[t][0], W[t][1], W[t][2].astype(np.float32)) for t in "KQV"}
    r = L.pipeline_ref_f32(mask, W, x.astype(np.float32), n, kdim)
    out32[f"case{i}_params"] = np.array([seed, m, kdim, n, band, 0], dtype=np.int64)
    out32[f"case{i}_fill"] = fill
    out32[f"case{i}_mask_values"] = mv
    for k, v in r.items():
        out32[f"case{i}_{k}"] = v
    out32["ncases"] = i + 1
np.savez_compressed(os.path.join(HERE, "pipeline_reference_f32.npz"), **out32)
print("wrote", out32["ncases"], "fp32 cases")
