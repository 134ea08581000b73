"""Regenerates tests/golden/*.  Run in the build container (needs /root/reference compiled
into oracle/_ref by `make -C oracle ref`):   python tests/golden/make_golden.py

Everything written here comes from the REFERENCE's own code:
  * matrices            <- artificial_matrix_generation()          (oracle/_ref/libamg_ref.so)
  * SpMM outputs        <- csr_to_format() + Matrix_Format::spmm() (oracle/_ref/libspmm_ref.so,
                           i.e. spmm_kernel_csr.cpp compiled unmodified)
  * partition vectors   <- loop_partitioner_balance_prefix_sums / _iterations macros and
                           MergePathSearch (libspmm_ref.so / libmerge_ref.so)
The FNV-1a-64 hashes pin the generator output (row_ptr / col_ind / values raw bytes).
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.abspath(os.path.join(HERE, "..", "..")))
from oracle import loader as L  # noqa: E402

# generator argument strings (harness order: rows cols avg std distr placement bw skew neigh crs seed)
PIN_CASES = [
    "100000 100000 10 2 normal random 1 0 0 0 14",                 # BASELINE configs[0]
    "200000 200000 20 10 gamma random 0.3 1000 0.5 0.5 14",        # skewed, SURVEY 8(c) row 2
    "50000 50000 40 5 normal diagonal 0.002 0 1.9 0.9 14",         # banded, SURVEY 8(c) row 3
    "3000 3000 8 3 normal random 0.5 0 0.1 0.1 7",
    "2000 5000 5 4 gamma random 1 100 0 0 3",                      # rectangular, empty rows, one long row
]
SPMM_CASES = [  # (args, k) small enough to commit the reference output
    ("3000 3000 8 3 normal random 0.5 0 0.1 0.1 7", 1),
    ("3000 3000 8 3 normal random 0.5 0 0.1 0.1 7", 8),
    ("3000 3000 8 3 normal random 0.5 0 0.1 0.1 7", 32),
    ("2000 5000 5 4 gamma random 1 100 0 0 3", 32),
    ("2000 5000 5 4 gamma random 1 100 0 0 3", 5),
]


def dense_b(c, k, seed):
    """Seeded non-constant B in the oracle layout x[n*c + j] (a fixed 64-bit LCG, no library RNG)."""
    n = c * k
    idx = np.arange(n, dtype=np.uint64)
    s = (idx + np.uint64(seed)) * np.uint64(6364136223846793005) + np.uint64(1442695040888963407)
    s ^= s >> np.uint64(33)
    s *= np.uint64(0xff51afd7ed558ccd)
    s ^= s >> np.uint64(33)
    return (s >> np.uint64(11)).astype(np.float64) / float(1 << 53)


def main():
    pins = {}
    for args in PIN_CASES:
        ia, ja, a, m, c = L.generate(args)
        nnz = len(ja)
        P = 8
        nnz_split = np.zeros(P + 1, np.int64); L.ref().ref_balance_prefix_sums(ia, m, nnz, P, nnz_split)
        rows_split = np.zeros(2 * P, np.int64); L.ref().ref_balance_iterations(0, m, P, rows_split)
        merge = np.zeros(2 * (P + 1), np.int64); L.merge_ref().ref_merge_partition(ia, m, nnz, P, merge)
        nnz_split3 = np.zeros(4, np.int64); L.ref().ref_balance_prefix_sums(ia, m, nnz, 3, nnz_split3)
        pins[args] = {
            "m": m, "c": c, "nnz": nnz, "max_row": int(np.diff(ia).max()),
            "fnv_row_ptr": L.fnv(ia), "fnv_col_ind": L.fnv(ja), "fnv_values": L.fnv(a),
            "nnz_split_P8": nnz_split.tolist(), "nnz_split_P3": nnz_split3.tolist(),
            "rows_split_P8": rows_split.tolist(), "merge_P8": merge.tolist(),
        }
        print(args, "->", nnz, pins[args]["fnv_row_ptr"])
    with open(os.path.join(HERE, "generator_pins.json"), "w") as f:
        json.dump(pins, f, indent=1)

    out = {}
    for i, (args, k) in enumerate(SPMM_CASES):
        ia, ja, a, m, c = L.generate(args)
        x = dense_b(c, k, seed=1000 + i)
        y = L.spmm_ref(ia, ja, a, m, c, x, k)
        out[f"case{i}_args"] = np.array(args)
        out[f"case{i}_k"] = np.array(k)
        out[f"case{i}_seed"] = np.array(1000 + i)
        out[f"case{i}_y"] = y
        out[f"case{i}_ia"], out[f"case{i}_ja"], out[f"case{i}_a"] = ia, ja, a   # so the case runs without the generator
        out[f"case{i}_fnv_y"] = np.array(L.fnv(y))
    out["ncases"] = np.array(len(SPMM_CASES))
    np.savez_compressed(os.path.join(HERE, "spmm_reference_outputs.npz"), **out)
    print("wrote", len(SPMM_CASES), "reference SpMM outputs")

    # COO -> CSR: output of the reference's coo_to_csr (csr_gen.c:162-217) on a shuffled COO of a
    # small generated matrix, plain and transposed.
    ia, ja, a, m, c = L.generate("3000 2500 12 6 gamma random 0.5 50 0.3 0.3 14")
    R = np.repeat(np.arange(m, dtype=np.int32), np.diff(ia))
    perm = np.random.default_rng(7).permutation(len(ja))
    R, C, V = R[perm].copy(), ja[perm].copy(), a[perm].copy()
    rp, ci, va = L.coo_to_csr_ref(R, C, V, m, c, 1, 0)
    assert np.array_equal(rp, ia) and np.array_equal(ci, ja) and np.array_equal(va, a)
    rpt, cit, vat = L.coo_to_csr_ref(R, C, V, m, c, 1, 1)
    np.savez_compressed(os.path.join(HERE, "coo_to_csr_reference.npz"), R=R, C=C, V=V, m=m, n=c,
                        row_ptr=rp, col_idx=ci, values=va, row_ptr_t=rpt, col_idx_t=cit, values_t=vat)
    print("wrote coo_to_csr reference vector, nnz =", len(R))


if __name__ == "__main__":
    main()
