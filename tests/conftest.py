import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def pkg():
    """The product package (spmm-research_b200/) -- raises if libspmm_b200.so is not built."""
    import __graft_entry__ as g
    p = g.load_package()
    p.load_library()
    return p


@pytest.fixture(scope="session")
def L():
    """The checker: oracle restatement + (when built) the compiled reference."""
    from oracle import loader
    loader.oracle()
    return loader


@pytest.fixture(scope="session")
def pins():
    with open(os.path.join(GOLDEN, "generator_pins.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def golden_cases():
    z = np.load(os.path.join(GOLDEN, "spmm_reference_outputs.npz"))
    cases = []
    for i in range(int(z["ncases"])):
        cases.append(dict(args=str(z[f"case{i}_args"]), k=int(z[f"case{i}_k"]), seed=int(z[f"case{i}_seed"]),
                          y=z[f"case{i}_y"], ia=z[f"case{i}_ia"], ja=z[f"case{i}_ja"], a=z[f"case{i}_a"],
                          fnv_y=str(z[f"case{i}_fnv_y"])))
    return cases


def dense_b(c, k, seed):
    """Same seeded B as tests/golden/make_golden.py (oracle layout x[n*c + j])."""
    n = c * k
    idx = np.arange(n, dtype=np.uint64)
    s = (idx + np.uint64(seed)) * np.uint64(6364136223846793005) + np.uint64(1442695040888963407)
    s ^= s >> np.uint64(33)
    s *= np.uint64(0xff51afd7ed558ccd)
    s ^= s >> np.uint64(33)
    return (s >> np.uint64(11)).astype(np.float64) / float(1 << 53)


def random_csr(m, c, avg, seed, empty_frac=0.0, long_row=0):
    """Generator-free CSR with sorted unique columns (for edge cases the generator cannot make)."""
    rng = np.random.default_rng(seed)
    deg = rng.poisson(avg, m).astype(np.int64)
    deg = np.minimum(deg, c)
    if empty_frac > 0:
        deg[rng.random(m) < empty_frac] = 0
    if long_row and m > 0:
        deg[m // 3] = min(long_row, c)
    ia = np.zeros(m + 1, np.int32)
    ia[1:] = np.cumsum(deg)
    cols = [np.sort(rng.choice(c, d, replace=False)) for d in deg]
    ja = (np.concatenate(cols) if len(cols) and ia[-1] > 0 else np.zeros(0)).astype(np.int32)
    a = rng.random(int(ia[-1])) - 0.25
    return ia, ja, a


def max_rel_err(y, y_ref):
    """max_i |y - y_ref| / max(|y_ref|, tiny): the SURVEY 8(c) parity measure."""
    if y.size == 0:
        return 0.0
    scale = np.maximum(np.abs(y_ref), 1e-300)
    return float(np.max(np.abs(y - y_ref) / scale))


FP64_TOL = 1e-12   # north_star: "within a stated relative tolerance (e.g. 1e-12 for fp64)"
