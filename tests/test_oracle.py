"""CPU tests: pin the oracle (our C restatement) against the golden vectors and, when the
compiled reference is present (oracle/_ref, build container), against the reference itself."""
import numpy as np
import pytest

from conftest import dense_b, max_rel_err, random_csr, FP64_TOL

# SURVEY.md 8(c) integer pins, produced by the surveyor with the reference's own macros
SURVEY_NNZ_SPLIT_P8 = [0, 12538, 25066, 37559, 50031, 62516, 75066, 87540, 100000]
SURVEY_MERGE_P8 = [(0, 0), (12534, 124901), (25059, 249811), (37554, 374751), (50028, 499712),
                   (62514, 624661), (75059, 749551), (87536, 874509), (100000, 999475)]
SURVEY_NNZ_SPLIT_P4_M2 = [0, 49249, 99499, 149750, 200000]
SURVEY_MERGE_P4_M2 = [(0, 0), (49285, 1005721), (99523, 2010489), (149761, 3015257), (200000, 4020021)]
CFG1 = "100000 100000 10 2 normal random 1 0 0 0 14"
CFG_SKEW = "200000 200000 20 10 gamma random 0.3 1000 0.5 0.5 14"


def test_golden_spmm_oracle_matches_reference_outputs(L, golden_cases):
    """Oracle vs outputs of the reference's compute_csr committed under tests/golden."""
    for g in golden_cases:
        m, c = len(g["ia"]) - 1, int(g["args"].split()[1])
        x = dense_b(c, g["k"], g["seed"])
        y = L.spmm_oracle(g["ia"], g["ja"], g["a"], m, c, x, g["k"])
        assert max_rel_err(y, g["y"]) <= 1e-15, g["args"]


def test_golden_merge_oracle_matches_reference_outputs(L, golden_cases):
    for g in golden_cases:
        m, c = len(g["ia"]) - 1, int(g["args"].split()[1])
        x = dense_b(c, g["k"], g["seed"])
        for P in (1, 7, 64):
            y = np.empty(m * g["k"])
            L.oracle().oracle_spmm_merge(g["ia"], g["ja"], g["a"], m, c, x, y, g["k"], P)
            assert max_rel_err(y, g["y"]) <= FP64_TOL


def test_partition_pins_small(L, pins, golden_cases):
    """Oracle partitioners vs vectors the reference macros produced (generator_pins.json),
    on the matrices committed in the fixture (no generator needed)."""
    for g in golden_cases[:4:3]:
        p = pins[g["args"]]
        ia = g["ia"]; m = len(ia) - 1; nnz = int(ia[-1])
        assert (m, nnz) == (p["m"], p["nnz"])
        out = np.zeros(9, np.int64); L.oracle().oracle_balance_prefix_sums(ia, m, nnz, 8, out)
        assert out.tolist() == p["nnz_split_P8"]
        out = np.zeros(4, np.int64); L.oracle().oracle_balance_prefix_sums(ia, m, nnz, 3, out)
        assert out.tolist() == p["nnz_split_P3"]
        out = np.zeros(16, np.int64); L.oracle().oracle_balance_iterations(0, m, 8, out)
        assert out.tolist() == p["rows_split_P8"]
        out = np.zeros(18, np.int64); L.oracle().oracle_merge_partition(ia, m, nnz, 8, out)
        assert out.tolist() == p["merge_P8"]
        assert L.fnv(ia) == p["fnv_row_ptr"] and L.fnv(g["ja"]) == p["fnv_col_ind"] and L.fnv(g["a"]) == p["fnv_values"]


def _need_ref(L):
    if not L.have_ref():
        pytest.skip("oracle/_ref not built (needs /root/reference)")


def test_generator_pins(L, pins):
    """The compiled reference generator reproduces the committed hashes (glibc random_r) and
    the oracle partitioners reproduce the reference's vectors on every pinned matrix."""
    _need_ref(L)
    for args, p in pins.items():
        ia, ja, a, m, c = L.generate(args)
        nnz = len(ja)
        assert (m, c, nnz, int(np.diff(ia).max())) == (p["m"], p["c"], p["nnz"], p["max_row"]), args
        assert (L.fnv(ia), L.fnv(ja), L.fnv(a)) == (p["fnv_row_ptr"], p["fnv_col_ind"], p["fnv_values"]), args
        out = np.zeros(9, np.int64); L.oracle().oracle_balance_prefix_sums(ia, m, nnz, 8, out)
        assert out.tolist() == p["nnz_split_P8"], args
        out = np.zeros(18, np.int64); L.oracle().oracle_merge_partition(ia, m, nnz, 8, out)
        assert out.tolist() == p["merge_P8"], args
        # CSR invariants the kernels rely on: sorted, unique columns inside a row
        d = np.diff(ja.astype(np.int64))
        row_starts = ia[1:-1][ia[1:-1] < nnz]
        ok = d > 0
        ok[row_starts[row_starts > 0] - 1] = True
        assert ok.all(), args


def test_survey_pins(L):
    """partition pins on configs[0] and on the skewed 200k matrix."""
    _need_ref(L)
    ia, ja, a, m, c = L.generate(CFG1)
    nnz = len(ja)
    assert nnz == 999475 and int(np.diff(ia).max()) == 19
    out = np.zeros(9, np.int64); L.oracle().oracle_balance_prefix_sums(ia, m, nnz, 8, out)
    assert out.tolist() == SURVEY_NNZ_SPLIT_P8
    assert np.diff(ia[out]).tolist() == [124939, 124930, 124933, 124940, 124930, 124936, 124930, 124937]
    out = np.zeros(16, np.int64); L.oracle().oracle_balance_iterations(0, m, 8, out)
    assert out.reshape(-1, 2)[:, 0].tolist() == [12500 * t for t in range(8)]
    out = np.zeros(18, np.int64); L.oracle().oracle_merge_partition(ia, m, nnz, 8, out)
    assert [tuple(r) for r in out.reshape(-1, 2).tolist()] == SURVEY_MERGE_P8
    ia, ja, a, m, c = L.generate(CFG_SKEW)
    nnz = len(ja)
    assert nnz == 4020021 and int(np.diff(ia).max()) == 20040
    out = np.zeros(5, np.int64); L.oracle().oracle_balance_prefix_sums(ia, m, nnz, 4, out)
    assert out.tolist() == SURVEY_NNZ_SPLIT_P4_M2
    out = np.zeros(10, np.int64); L.oracle().oracle_merge_partition(ia, m, nnz, 4, out)
    assert [tuple(r) for r in out.reshape(-1, 2).tolist()] == SURVEY_MERGE_P4_M2


@pytest.mark.parametrize("seed", range(6))
def test_oracle_vs_compiled_reference_random(L, seed):
    """Oracle restatement == the reference's own code on random inputs incl. ragged/empty rows."""
    _need_ref(L)
    rng = np.random.default_rng(seed)
    m, c = int(rng.integers(1, 3000)), int(rng.integers(1, 3000))
    ia, ja, a = random_csr(m, c, avg=float(rng.uniform(0.3, 12)), seed=seed, empty_frac=0.2,
                           long_row=int(rng.integers(0, 2)) * min(c, 900))
    nnz = len(ja)
    k = int(rng.choice([1, 2, 5, 8, 32]))
    x = dense_b(c, k, seed)
    assert np.array_equal(L.spmm_oracle(ia, ja, a, m, c, x, k), L.spmm_ref(ia, ja, a, m, c, x, k))
    for P in (1, 2, 3, 8, 61, 1000):
        o = np.zeros(P + 1, np.int64); L.oracle().oracle_balance_prefix_sums(ia, m, nnz, P, o)
        r = np.zeros(P + 1, np.int64); L.ref().ref_balance_prefix_sums(ia, m, nnz, P, r)
        assert np.array_equal(o, r), (seed, P)
        o = np.zeros(2 * P, np.int64); L.oracle().oracle_balance_iterations(0, m, P, o)
        r = np.zeros(2 * P, np.int64); L.ref().ref_balance_iterations(0, m, P, r)
        assert np.array_equal(o, r), (seed, P)
        o = np.zeros(2 * P + 2, np.int64); L.oracle().oracle_merge_partition(ia, m, nnz, P, o)
        r = np.zeros(2 * P + 2, np.int64); L.merge_ref().ref_merge_partition(ia, m, nnz, P, r)
        assert np.array_equal(o, r), (seed, P)
    # k=1: the reference's CPU merge-path SpMV against the oracle's k-wide generalisation
    x1 = dense_b(c, 1, seed + 99)
    y_ref = np.zeros(m); L.merge_ref().ref_spmv_merge(ia, ja, a, m, nnz, x1, y_ref, 5)
    y_or = np.zeros(m); L.oracle().oracle_spmm_merge(ia, ja, a, m, c, x1, y_or, 1, 5)
    assert np.allclose(y_ref, y_or, rtol=1e-12, atol=1e-13)   # FMA contraction differs between the two TUs


def test_nnz_split_worker_ends_match_next_starts(L):
    """The macro returns (start, end) per worker; ends equal the next worker's start, last = m."""
    _need_ref(L)
    ia, ja, a = random_csr(500, 400, 6.0, 3, empty_frac=0.3)
    m, nnz, P = 500, len(ja), 7
    pairs = np.zeros(2 * P, np.int64); L.ref().ref_balance_prefix_sums_pairs(ia, m, nnz, P, pairs)
    pairs = pairs.reshape(-1, 2)
    o = np.zeros(P + 1, np.int64); L.oracle().oracle_balance_prefix_sums(ia, m, nnz, P, o)
    assert pairs[:, 0].tolist() == o[:-1].tolist() and pairs[:, 1].tolist() == o[1:].tolist()
