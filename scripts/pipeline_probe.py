"""Times the sparse-attention pipeline (N4) stage by stage on one B200 and against the CPU restatement.
   python scripts/pipeline_probe.py [length] [n] [fill] [band]
Prints one markdown table row per stage: kernel ms (CUDA events inside the library), algorithmic bytes, GB/s."""
import os
import sys
import time

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as g                      # noqa: E402
from oracle import loader as L                   # noqa: E402  (checker + CPU baseline only)
from pipeline_util import pipeline_case          # noqa: E402

length = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
n = int(sys.argv[2]) if len(sys.argv) > 2 else 64
fill = float(sys.argv[3]) if len(sys.argv) > 3 else 0.05
band = int(sys.argv[4]) if len(sys.argv) > 4 else 16
kdim = length
pkg = g.load_package()
t0 = time.time()
mask, W, x = pipeline_case(3, length, kdim, n, band, fill, avg_w=64.0)
mia, mja, ma, m = mask
nnz = len(mja)
wnnz = {t: len(W[t][1]) for t in "KQV"}
print(f"# pipeline probe: length {length}, n {n}, mask nnz {nnz} (fill {fill}, band {band}), weights nnz {wnnz}, gen {time.time() - t0:.1f}s")

MF = pkg.csr_to_pipeline_format(mia, mja, ma, m, nnz, n)
for t in "KQV":
    MF.set_weight(t, m, kdim, *W[t])
yf = np.empty(m * n)
REP = 20


def best(fn):
    v = []
    for _ in range(REP):
        fn()
        v.append(MF.last_kernel_ms())
    return float(np.median(v[3:]))


rows = []
out = {t: np.empty(m * n) for t in "KQV"}
for t in "KQV":
    ms = best(lambda: MF.spmm(t, m, kdim, n, *W[t], x, None))
    byt = wnnz[t] * (12 + 8 * n) + 8 * m * n + 4 * m
    rows.append((f"spmm('{t}')", ms, byt))
for mode, name in ((0, "sddmm reference mode"), (1, "sddmm attention mode"), (2, "sddmm attention mode, ordered")):
    ms = best(lambda: MF.sddmm(None, mode))
    byt = (16 * n * m + 16 * nnz + 4 * m) if mode == 0 else (8 * n * m + nnz * (8 * n + 20) + 4 * m)
    rows.append((name, ms, byt))
ms = best(lambda: MF.softmax(None, rowwise=True))
rows.append(("softmax (row-wise)", ms, 16 * nnz))
MF.sddmm(None, 1)
ms = best(lambda: MF.softmax(None, rowwise=False))
rows.append(("softmax (global, 5 kernels)", ms, 32 * nnz))
MF.sddmm(None, 1)
ms = best(lambda: MF.final(yf))
rows.append(("spmm('final')", ms, nnz * (12 + 8 * n) + 8 * m * n + 4 * m))
for mode in (0, 1, 2):
    ms = best(lambda: MF.attention(x, yf, mode, 0))
    rows.append((f"attention(), mode {mode}, all kernels", ms, 0))
for fused in ("1", "0"):
    os.environ["SPMM_B200_ATTENTION_FUSED"] = fused
    ms = best(lambda: MF.attention(x, yf, 1, 2))
    rows.append((f"attention(), mode 1 + row softmax, {'ONE fused kernel' if fused == '1' else 'three kernels'} after the projections", ms, 0))
os.environ.pop("SPMM_B200_ATTENTION_FUSED")
print("| stage | kernel ms | algorithmic MB | GB/s |")
print("|---|---|---|---|")
for name, ms, byt in rows:
    print(f"| {name} | {ms:.4f} | {byt / 1e6:.1f} | {byt / ms / 1e6 if byt else 0:.0f} |")

# end to end through the host-pointer call (x up, y_final down), wall clock
ts = []
for _ in range(10):
    t0 = time.perf_counter()
    MF.attention(x, yf, 1, 0)
    ts.append(time.perf_counter() - t0)
e2e = float(np.median(ts[2:])) * 1e3
flops = 2.0 * n * (sum(wnnz.values()) + 2 * nnz)
print(f"attention() end to end (host x -> host y_final): {e2e:.3f} ms, {flops / e2e / 1e6:.1f} GFLOP/s "
      f"(2*n*(nnz_K + nnz_Q + nnz_V + 2*nnz_mask) flop)")

# CPU restatement (the reference's algorithm, all host threads), and parity of the full-size result
t0 = time.perf_counter()
MF.attention(x, yf, 2, 0)
r = L.pipeline_oracle(mask, W, x, n, 1, 0)
cpu = (time.perf_counter() - t0) * 1e3
print(f"CPU restatement, {L.oracle().oracle_num_threads()} threads: {cpu:.1f} ms, {flops / cpu / 1e6:.1f} GFLOP/s; "
      f"bit-identical y_final: {np.array_equal(r['y_final'], yf)}")
MF.close()
