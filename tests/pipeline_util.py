"""Inputs for the sparse-attention pipeline tests.

band_and_random_mask restates the SHAPE of the reference's mask generator
(pipeline_code_bench/sddmm_mask.h:17-81, band_and_random): a band |i - j| < band_size of ones plus uniformly
drawn lower-triangle entries until `fill` * length^2 ones are placed, converted row by row to CSR with value 1.0
(dense_to_csr, sddmm_mask.h:273-296).  The reference seeds rand() from time(NULL), so its masks are not
reproducible; here the draws come from a seeded numpy generator.  Test input only, not product code."""
import numpy as np

from conftest import random_csr


def band_and_random_mask(length, band_size, fill, seed):
    rng = np.random.default_rng(seed)
    dense = np.zeros((length, length), dtype=bool)
    i = np.arange(length)
    for off in range(-(band_size - 1), band_size):
        j = i + off
        ok = (j >= 0) & (j < length)
        dense[i[ok], j[ok]] = True
    want = int(fill * length * length)
    placed = int(dense.sum())
    while placed < want:
        r = rng.integers(0, length, size=2 * (want - placed) + 16)
        c = (rng.random(len(r)) * (r + 1)).astype(np.int64)          # col = rand() % (row + 1)
        flat = r * length + c
        flat = flat[~dense.ravel()[flat]]
        _, first = np.unique(flat, return_index=True)                # keep draw order, drop repeats
        flat = flat[np.sort(first)][:want - placed]
        dense.ravel()[flat] = True
        placed += len(flat)
    ia = np.zeros(length + 1, np.int32)
    ia[1:] = np.cumsum(dense.sum(axis=1))
    ja = np.nonzero(dense)[1].astype(np.int32)
    return ia, ja, np.ones(len(ja), np.float64)


def pipeline_case(seed, m, kdim, n, band_size=4, fill=0.1, avg_w=6.0, mask_values="ones", empty_frac=0.0):
    """mask (ia, ja, a, m), weights {'K','Q','V': (ia, ja, a)} (m x kdim CSR), x (kdim*n, row-major)."""
    mia, mja, ma = band_and_random_mask(m, band_size, fill, seed)
    rng = np.random.default_rng(seed + 1000)
    if mask_values != "ones":
        ma = rng.random(len(mja)) + 0.5
    W = {}
    for i, t in enumerate("KQV"):
        W[t] = random_csr(m, kdim, avg_w, seed * 7 + i, empty_frac=empty_frac)
    x = rng.standard_normal(kdim * n)
    return (mia, mja, ma, m), W, x
