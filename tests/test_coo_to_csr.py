"""COO -> CSR: the oracle restatement is pinned against the reference's own
coo_to_csr (csr_gen.c:162-217, compiled into oracle/_ref/libcoo_ref.so) and a committed golden
vector; the device conversion (spmm_b200_coo_to_csr) must match the oracle exactly."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, random_csr


def coo_from_csr(ia, ja, a, seed):
    """Shuffled COO triplets of a CSR matrix (unique (row, col) pairs)."""
    m = len(ia) - 1
    R = np.repeat(np.arange(m, dtype=np.int32), np.diff(ia))
    perm = np.random.default_rng(seed).permutation(len(ja))
    return R[perm].copy(), ja[perm].astype(np.int32).copy(), a[perm].copy()


CASES = [  # m, c, avg, empty_frac, long_row
    (1, 7, 3.0, 0.0, 0),
    (300, 200, 5.0, 0.3, 0),
    (2000, 1500, 20.0, 0.05, 1400),      # one row longer than n/5 -> reference takes its bucket-sort path
    (64, 6000, 60.0, 0.0, 5000),         # row longer than a CTA's shared-memory sort
]


@pytest.mark.parametrize("case", range(len(CASES)))
def test_oracle_matches_reference_coo_to_csr(L, case):
    if not L.have_coo_ref():
        pytest.skip("oracle/_ref/libcoo_ref.so not built")
    m, c, avg, ef, lr = CASES[case]
    ia, ja, a = random_csr(m, c, avg, seed=case, empty_frac=ef, long_row=lr)
    R, C, V = coo_from_csr(ia, ja, a, case)
    for transpose in (0, 1):
        o = L.coo_to_csr_oracle(R, C, V, m, c, 1, transpose)
        r = L.coo_to_csr_ref(R, C, V, m, c, 1, transpose)
        assert np.array_equal(o[0], r[0]) and np.array_equal(o[1], r[1]) and np.array_equal(o[2], r[2])
    o = L.coo_to_csr_oracle(R, C, V, m, c, 1, 0)
    assert np.array_equal(o[0], ia) and np.array_equal(o[1], ja) and np.array_equal(o[2], a)


def test_oracle_coo_golden(L):
    """Committed vector produced by the reference (tests/golden/make_golden.py)."""
    z = np.load(os.path.join(GOLDEN, "coo_to_csr_reference.npz"))
    o = L.coo_to_csr_oracle(z["R"], z["C"], z["V"], int(z["m"]), int(z["n"]), 1, 0)
    assert np.array_equal(o[0], z["row_ptr"]) and np.array_equal(o[1], z["col_idx"]) and np.array_equal(o[2], z["values"])
    ot = L.coo_to_csr_oracle(z["R"], z["C"], z["V"], int(z["m"]), int(z["n"]), 1, 1)
    assert np.array_equal(ot[0], z["row_ptr_t"]) and np.array_equal(ot[1], z["col_idx_t"]) and np.array_equal(ot[2], z["values_t"])


def test_oracle_coo_empty_and_duplicates(L):
    o = L.coo_to_csr_oracle(np.zeros(0, np.int32), np.zeros(0, np.int32), np.zeros(0), 5, 4, 1, 0)
    assert o[0].tolist() == [0] * 6 and len(o[1]) == 0
    # duplicates: ordered by (column, input position); unsorted mode keeps input order per row
    R = np.array([1, 0, 1, 1, 0], np.int32); C = np.array([2, 3, 2, 0, 3], np.int32); V = np.array([1., 2., 3., 4., 5.])
    o = L.coo_to_csr_oracle(R, C, V, 2, 4, 1, 0)
    assert o[0].tolist() == [0, 2, 5] and o[1].tolist() == [3, 3, 0, 2, 2] and o[2].tolist() == [2., 5., 4., 1., 3.]
    u = L.coo_to_csr_oracle(R, C, V, 2, 4, 0, 0)
    assert u[1].tolist() == [3, 3, 2, 2, 0] and u[2].tolist() == [2., 5., 1., 3., 4.]


@pytest.mark.gpu
@pytest.mark.parametrize("case", range(len(CASES)))
def test_device_coo_to_csr_bit_exact(pkg, L, case):
    m, c, avg, ef, lr = CASES[case]
    ia, ja, a = random_csr(m, c, avg, seed=case, empty_frac=ef, long_row=lr)
    R, C, V = coo_from_csr(ia, ja, a, case)
    nnz = len(R)
    for transpose in (0, 1):
        for sort_columns in (1, 0):
            o = L.coo_to_csr_oracle(R, C, V, m, c, sort_columns, transpose)
            d = pkg.coo_to_csr(R, C, V, m, c, nnz, sort_columns, transpose)
            assert np.array_equal(d[0], o[0]), (case, transpose, sort_columns)
            assert np.array_equal(d[1], o[1]), (case, transpose, sort_columns)
            assert np.array_equal(d[2], o[2]), (case, transpose, sort_columns)
    d = pkg.coo_to_csr(R, C, None, m, c, nnz)             # pattern only
    assert d[2] is None and np.array_equal(d[1], ja)


@pytest.mark.gpu
def test_device_coo_edge_cases(pkg, L):
    e = np.zeros(0, np.int32)
    d = pkg.coo_to_csr(e, e, np.zeros(0), 5, 4, 0)
    assert d[0].tolist() == [0] * 6
    R = np.array([1, 0, 1, 1, 0], np.int32); C = np.array([2, 3, 2, 0, 3], np.int32); V = np.array([1., 2., 3., 4., 5.])
    d = pkg.coo_to_csr(R, C, V, 2, 4, 5)
    assert d[1].tolist() == [3, 3, 0, 2, 2] and d[2].tolist() == [2., 5., 4., 1., 3.]
    with pytest.raises(pkg.SpmmB200Error):
        pkg.coo_to_csr(np.array([7], np.int32), np.array([0], np.int32), np.array([1.]), 2, 4, 1)


@pytest.mark.gpu
def test_device_coo_generated_matrix_roundtrip(pkg, L, pins):
    """A generator matrix shuffled into COO converts back to the identical CSR (pinned hashes)."""
    if not L.have_ref():
        pytest.skip("reference generator not built")
    args = "200000 200000 20 10 gamma random 0.3 1000 0.5 0.5 14"
    ia, ja, a, m, c = L.generate(args)
    R, C, V = coo_from_csr(ia, ja, a, 1)
    d = pkg.coo_to_csr(R, C, V, m, c, len(R))
    assert L.fnv(d[0]) == pins[args]["fnv_row_ptr"] and L.fnv(d[1]) == pins[args]["fnv_col_ind"]
    assert L.fnv(d[2]) == pins[args]["fnv_values"]
