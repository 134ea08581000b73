""""perfect nnz balance with row carry" (spmv_kernel_csr.cpp:68-80 set-up, :625-680 compute).
The partition must be bit-exact: oracle restatement == the reference's own TU built as Custom_CSR_PBV
(oracle/_ref/libpbv_ref.so, when present) == committed vectors the reference produced (tests/golden/pbv_pins.json,
made by tests/golden/make_pbv_golden.py) == spmm_b200_partition(SPMM_B200_PART_NNZ_CARRY) on the host and on the device."""
import json
import os

import numpy as np
import pytest

from conftest import GOLDEN, FP64_TOL, dense_b, max_rel_err, random_csr


@pytest.fixture(scope="module")
def pbv_pins():
    with open(os.path.join(GOLDEN, "pbv_pins.json")) as f:
        return json.load(f)


def _oracle_part(L, ia, m, nnz, P):
    o = np.zeros(4 * P, np.int64)
    L.oracle().oracle_nnz_carry_partition(ia, m, nnz, P, o)
    return o


def test_oracle_matches_reference_vectors(L, pbv_pins, golden_cases):
    """On the matrices committed in the SpMM fixture (no generator needed): partition and SpMV result hash."""
    seen = 0
    for g in golden_cases:
        if g["args"] not in pbv_pins:
            continue
        p = pbv_pins[g["args"]]
        ia, ja, a = g["ia"], g["ja"], g["a"]
        m, nnz, c = len(ia) - 1, int(ia[-1]), int(g["args"].split()[1])
        assert (m, nnz) == (p["m"], p["nnz"])
        x = dense_b(c, 1, 77)
        for P in (3, 8, 148):
            part = _oracle_part(L, ia, m, nnz, P)
            assert (part.tolist() if P <= 8 else L.fnv(part)) == p[f"P{P}"]
            y = np.full(m, np.nan)
            L.oracle().oracle_spmv_nnz_carry(ia, ja, a, m, nnz, P, x, y)
            assert L.fnv(y) == p[f"fnv_y_P{P}"]            # same bits as the reference's kernel
            assert max_rel_err(y, L.spmm_oracle(ia, ja, a, m, c, x, 1)) <= FP64_TOL
        seen += 1
    assert seen >= 2


def test_host_partition_matches_oracle_on_ragged_inputs(pkg, L):
    for seed in range(8):
        m = 1 + 613 * seed
        ia, ja, a = random_csr(m, 400, 3.0, seed, empty_frac=0.5, long_row=(seed % 2) * 350)
        nnz = len(ja)
        for P in (1, 2, 5, 33, 148, 3000):
            o = _oracle_part(L, ia, m, nnz, P)
            assert np.array_equal(pkg.partition(ia, m, nnz, P, pkg.PART_NNZ_CARRY, host=True).ravel(), o)
    # degenerate shapes: no rows, no nonzeros
    z = np.zeros(1, np.int32)
    assert pkg.partition(z, 0, 0, 4, pkg.PART_NNZ_CARRY, host=True).tolist() == _oracle_part(L, z, 0, 0, 4).reshape(-1, 4).tolist()
    z = np.zeros(8, np.int32)
    assert pkg.partition(z, 7, 0, 3, pkg.PART_NNZ_CARRY, host=True).tolist() == _oracle_part(L, z, 7, 0, 3).reshape(-1, 4).tolist()


def test_oracle_matches_compiled_reference(L):
    if not L.have_pbv_ref():
        pytest.skip("oracle/_ref/libpbv_ref.so not built (needs /root/reference)")
    for seed in range(6):
        m = 50 + 997 * seed
        ia, ja, a = random_csr(m, 500, 5.0, 100 + seed, empty_frac=0.3, long_row=(seed % 3 == 0) * 450)
        nnz = len(ja)
        x = np.random.default_rng(seed).random(500)
        for P in (1, 4, 24, 200):
            r = np.zeros(4 * P, np.int64)
            L.pbv_ref().ref_pbv_partition(ia, ja, a, m, 500, nnz, P, r)
            assert np.array_equal(_oracle_part(L, ia, m, nnz, P), r), (seed, P)
            y_ref = np.full(m, np.nan); L.pbv_ref().ref_pbv_spmv(ia, ja, a, m, 500, nnz, P, x, y_ref)
            y = np.full(m, np.nan); L.oracle().oracle_spmv_nnz_carry(ia, ja, a, m, nnz, P, x, y)
            assert np.array_equal(y, y_ref), (seed, P)


@pytest.mark.gpu
def test_device_partition_bit_exact(pkg, L, pbv_pins, golden_cases):
    for g in golden_cases:
        if g["args"] not in pbv_pins:
            continue
        p = pbv_pins[g["args"]]
        ia = g["ia"]; m = len(ia) - 1; nnz = int(ia[-1])
        for P in (3, 8):
            assert pkg.partition(ia, m, nnz, P, pkg.PART_NNZ_CARRY).ravel().tolist() == p[f"P{P}"]
        assert L.fnv(pkg.partition(ia, m, nnz, 148, pkg.PART_NNZ_CARRY).ravel()) == p["P148"]
    for seed in range(5):
        m = 1 + 977 * seed
        ia, ja, a = random_csr(m, 500, 4.0, seed, empty_frac=0.4, long_row=(seed % 2) * 400)
        nnz = len(ja)
        for P in (1, 2, 7, 148, 5000):
            assert np.array_equal(pkg.partition(ia, m, nnz, P, pkg.PART_NNZ_CARRY).ravel(), _oracle_part(L, ia, m, nnz, P))


@pytest.mark.gpu
def test_generated_matrices_device_partition(pkg, L, pbv_pins):
    """The pinned generator matrices, regenerated on the GPU box by the reference's generator."""
    if not L.have_ref():
        pytest.skip("reference generator not built")
    for args, p in pbv_pins.items():
        ia, ja, a, m, c = L.generate(args)
        nnz = len(ja)
        assert pkg.partition(ia, m, nnz, 8, pkg.PART_NNZ_CARRY).ravel().tolist() == p["P8"]
        # the k = 1 merge-path kernel is this repo's row-carry SpMV: same result within tolerance
        x = dense_b(c, 1, 77)
        y_ref = np.full(m, np.nan); L.oracle().oracle_spmv_nnz_carry(ia, ja, a, m, nnz, 8, x, y_ref)
        assert L.fnv(y_ref) == p["fnv_y_P8"]
        mf = pkg.csr_to_format(ia, ja, a, m, c, nnz, 1, variant=pkg.MERGE)
        y = np.full(m, np.nan); mf.spmm(x, y, 1); mf.close()
        assert max_rel_err(y, y_ref) <= FP64_TOL
