"""GPU tests of the row-band host pipeline (spmm_b200.cu run_rowband): on a banded matrix spmm_b200_run walks down the
rows -- B rows up, kernel on a row window, C rows down, block by block -- instead of slicing the k columns.  The result
must be the same bits as the k-slice pipeline and within 1e-12 of the oracle, whatever the block count, the layout of B,
the kind of host memory, or how un-banded the matrix is (the window then grows in both directions)."""
import numpy as np
import pytest

from conftest import FP64_TOL, dense_b, max_rel_err

pytestmark = pytest.mark.gpu


def banded_csr(m, c, avg, half_band, seed, empty_rows=(), wrap_first=False):
    """Sorted unique columns within +-half_band of the scaled diagonal; optional empty row ranges and a first row
    that reaches the last columns (an un-banded outlier)."""
    rng = np.random.default_rng(seed)
    deg = np.minimum(rng.poisson(avg, m).astype(np.int64), 2 * half_band)
    for r0, r1 in empty_rows:
        deg[r0:r1] = 0
    cols = []
    for i in range(m):
        centre = int(i * (c - 1) / max(1, m - 1))
        lo, hi = max(0, centre - half_band), min(c, centre + half_band + 1)
        d = int(min(deg[i], hi - lo))
        row = np.sort(rng.choice(np.arange(lo, hi), d, replace=False))
        if wrap_first and i == 0:
            row = np.unique(np.concatenate([row, [c - 1, c - 2]]))
        cols.append(row)
    ia = np.zeros(m + 1, np.int32)
    ia[1:] = np.cumsum([len(r) for r in cols])
    ja = np.concatenate(cols).astype(np.int32) if ia[-1] else np.zeros(0, np.int32)
    a = rng.random(int(ia[-1])) + 0.01          # positive: no cancellation, so the relative error is meaningful on every element
    return ia, ja, a


def _run(mf, x, k, m, pinned):
    if pinned:
        # one page-locked buffer per handle, kept alive with it (registered memory must not be freed before close())
        if getattr(mf, "_test_pinned", None) is None:
            mf._test_pinned = np.empty(m * k)
            mf.register_host(mf._test_pinned)
        y = mf._test_pinned
        y[:] = np.nan
        mf.spmm(x, y, k)
        return y.copy()
    y = np.full(m * k, np.nan)
    mf.spmm(x, y, k)
    return y


@pytest.mark.parametrize("k", [8, 16, 32, 64, 128])
def test_rowband_matches_kslice_and_oracle(pkg, L, k):
    m, c = 6000, 5000
    ia, ja, a = banded_csr(m, c, 24.0, 150, seed=k)
    x = dense_b(c, k, 3)
    y_ref = L.spmm_oracle(ia, ja, a, m, c, x, k)
    mf = pkg.csr_to_format(ia, ja, a, m, c, len(ja), k, variant=pkg.ROWSPLIT)
    mf.set_option("ROWBAND", 0); mf.set_option("PIPELINE_MIN_KB", 0)
    y0 = _run(mf, x, k, m, pinned=False)
    assert mf.last_rowband() == 0
    assert max_rel_err(y0, y_ref) <= FP64_TOL
    garbage = dense_b(c, k, 77) * 1e3 - 5e2
    for blocks in (1, 4, 7, 16, 64, -2):
        for pinned in (False, True):
            for direct in (1, 0):                # kernel stores into the mapped y / one contiguous copy per block
                mf.set_option("ROWBAND_DIRECT", direct)
                mf.set_option("ROWBAND", 0)      # a B left over from the previous run must not hide a missing upload:
                _run(mf, garbage, k, m, False)   # overwrite the device copies with another operand first
                mf.set_option("ROWBAND", blocks)
                y = _run(mf, x, k, m, pinned)
                assert mf.last_rowband() == (blocks if blocks > 0 else 7), (k, blocks)
                assert np.array_equal(y, y0), (k, blocks, pinned, direct)       # same kernel arithmetic, row by row
                assert len(mf.last_timeline()) == mf.last_rowband()
    # fetch_c after a staged (copy-engine) row-band run returns the same C
    mf.set_option("ROWBAND_DIRECT", 0); mf.set_option("ROWBAND", 8)
    y = _run(mf, x, k, m, pinned=True)
    y2 = np.full(m * k, np.nan); mf.fetch_c(y2, k)
    assert np.array_equal(y2, y0)
    mf.close()


def test_rowband_row_major_b(pkg, L):
    m, c, k = 5000, 7000, 32
    ia, ja, a = banded_csr(m, c, 30.0, 200, seed=5)
    x = dense_b(c, k, 9)
    y_ref = L.spmm_oracle(ia, ja, a, m, c, x, k)
    x_rm = np.ascontiguousarray(x.reshape(k, c).T).ravel()        # B[j*k + n]
    mf = pkg.csr_to_format(ia, ja, a, m, c, len(ja), k, variant=pkg.ROWSPLIT, b_layout=pkg.B_ROWMAJOR)
    mf.set_option("ROWBAND", 8)
    y = _run(mf, x_rm, k, m, pinned=True)
    assert mf.last_rowband() == 8
    assert max_rel_err(y, y_ref) <= FP64_TOL
    mf.close()


@pytest.mark.parametrize("case", ["random", "empty_blocks", "outlier_first_row", "anti_diagonal"])
def test_rowband_on_unbanded_and_ragged_matrices(pkg, L, case):
    from conftest import random_csr
    m, c, k = 4096, 3000, 32
    if case == "random":                       # every block touches every column: the first upload is all of B
        ia, ja, a = random_csr(m, c, 12.0, seed=21, empty_frac=0.1)
    elif case == "empty_blocks":               # whole row blocks without a nonzero, at the start and in the middle
        ia, ja, a = banded_csr(m, c, 20.0, 100, seed=22, empty_rows=[(0, 700), (2000, 2600)])
    elif case == "outlier_first_row":          # window jumps to the last column at once
        ia, ja, a = banded_csr(m, c, 20.0, 100, seed=23, wrap_first=True)
    else:                                      # band runs from the last columns to the first: the window grows downwards
        ia, ja, a = banded_csr(m, c, 20.0, 100, seed=24)
        ja = (c - 1 - ja).astype(np.int32)
        for i in range(m):                     # keep columns sorted inside a row
            ja[ia[i]:ia[i + 1]] = ja[ia[i]:ia[i + 1]][::-1]
            a[ia[i]:ia[i + 1]] = a[ia[i]:ia[i + 1]][::-1]
    x = dense_b(c, k, 4)
    y_ref = L.spmm_oracle(ia, ja, a, m, c, x, k)
    mf = pkg.csr_to_format(ia, ja, a, m, c, len(ja), k, variant=pkg.ROWSPLIT)
    for blocks in (8, 16):
        mf.set_option("ROWBAND", blocks)
        y = _run(mf, x, k, m, pinned=True)
        assert mf.last_rowband() == blocks
        assert max_rel_err(y, y_ref) <= FP64_TOL, (case, blocks)
    mf.close()


def banded_csr_fast(m, c, d, half_band, seed):
    """Vectorised: exactly d sorted distinct columns per row inside +-half_band of the scaled diagonal."""
    rng = np.random.default_rng(seed)
    gaps = rng.integers(1, max(2, 2 * half_band // d), (m, d))
    offs = np.cumsum(gaps, axis=1) - 1                                    # < 2 * half_band
    start = (np.arange(m, dtype=np.int64) * (c - 2 * half_band - 1) // max(1, m - 1))[:, None]
    ja = (start + offs).astype(np.int32).ravel()
    ia = (np.arange(m + 1, dtype=np.int64) * d).astype(np.int32)
    a = rng.random(m * d) + 0.01
    return ia, ja, a


def test_rowband_rule(pkg, L):
    """ROWBAND=-1 (default): a banded operand above the pipeline threshold takes the row-band pipeline, an un-banded one
    and every other variant keep the k-slice pipeline; small operands keep the single upload."""
    from conftest import random_csr
    m = c = 400000
    k = 32                                     # B = C = 102 MB
    ia, ja, a = banded_csr_fast(m, c, 8, 1000, seed=31)
    x = dense_b(c, k, 2)
    y_ref = L.spmm_oracle(ia, ja, a, m, c, x, k)
    mf = pkg.csr_to_format(ia, ja, a, m, c, len(ja), k, variant=pkg.ROWSPLIT)
    y = _run(mf, x, k, m, pinned=True)
    assert mf.last_rowband() in (8, 16)                                   # diagonal band: equal blocks
    assert max_rel_err(y, y_ref) <= FP64_TOL
    up, down = mf.last_transfer_bytes()
    assert down == m * k * 8 and 0.99 * c * k * 8 <= up <= c * k * 8      # every row of B is needed exactly once
    mf.close()
    mg = pkg.csr_to_format(ia, ja, a, m, c, len(ja), k, variant=pkg.MERGE)
    _run(mg, x, k, m, pinned=True)
    assert mg.last_rowband() == 0
    mg.close()
    # a column stripe (the reference generator's "random" placement with bw < 1): short upload, graded blocks
    js = (ja.astype(np.int64) % 20000 + 150000).astype(np.int32)
    js = np.sort(js.reshape(m, 8), axis=1)
    keep = np.ones_like(js, bool); keep[:, 1:] = js[:, 1:] != js[:, :-1]  # drop duplicate columns inside a row
    ias = np.zeros(m + 1, np.int32); ias[1:] = np.cumsum(keep.sum(axis=1))
    jas, a_s = js[keep], a.reshape(m, 8)[keep]
    mst = pkg.csr_to_format(ias, jas, a_s, m, c, len(jas), k, variant=pkg.ROWSPLIT)
    y = _run(mst, x, k, m, pinned=True)
    assert mst.last_rowband() == 7
    assert max_rel_err(y, L.spmm_oracle(ias, jas, a_s, m, c, x, k)) <= FP64_TOL
    up, down = mst.last_transfer_bytes()
    assert up == (int(jas.max()) - int(jas.min()) + 1) * k * 8 and down == m * k * 8
    mst.close()
    ia, ja, a = random_csr(m // 4, c, 8.0, seed=32)                       # un-banded: the first block needs all of B
    mr = pkg.csr_to_format(ia, ja, a, m // 4, c, len(ja), k, variant=pkg.ROWSPLIT)
    _run(mr, x, k, m // 4, pinned=True)
    assert mr.last_rowband() == 0
    mr.close()
    ia, ja, a = banded_csr(3000, 3000, 8.0, 50, seed=33)                  # small: one upload, one kernel, one download
    ms = pkg.csr_to_format(ia, ja, a, 3000, 3000, len(ja), k, variant=pkg.ROWSPLIT)
    _run(ms, dense_b(3000, k, 2), k, 3000, pinned=False)
    assert ms.last_rowband() == 0
    ms.close()


@pytest.mark.parametrize("layout", ["colmajor", "rowmajor"])
def test_upload_window(pkg, L, layout):
    """Only rows [first touched column, last touched column] of B are uploaded, on every host-buffer path; what lies
    outside may be anything (NaN here) -- the reference never reads it either."""
    m, c, k = 5000, 9000, 32
    lo, hi = 2100, 2900                                                    # the stripe of touched columns
    ia, ja, a = banded_csr(m, hi - lo, 18.0, 60, seed=41)
    ja = (ja + lo).astype(np.int32)
    x = dense_b(c, k, 6)
    y_ref = L.spmm_oracle(ia, ja, a, m, c, x, k)
    xp = x.reshape(k, c).copy()
    xp[:, :int(ja.min())] = np.nan; xp[:, int(ja.max()) + 1:] = np.nan
    span = int(ja.max()) - int(ja.min()) + 1
    if layout == "rowmajor":
        xp = np.ascontiguousarray(xp.T)
    xp = xp.ravel()
    mf = pkg.csr_to_format(ia, ja, a, m, c, len(ja), k, variant=pkg.ROWSPLIT,
                           b_layout=pkg.B_ROWMAJOR if layout == "rowmajor" else pkg.B_COLMAJOR)
    mf.set_option("UPLOAD_WINDOW", 0); mf.set_option("ROWBAND", 0)        # poison the device copies of B: NaN everywhere
    mf.spmm(np.full(c * k, np.nan), np.empty(m * k), k)
    mf.set_option("UPLOAD_WINDOW", 1)
    for name, opts in [("single", {"ROWBAND": 0}), ("kslice", {"ROWBAND": 0, "PIPELINE_MIN_KB": 0}),
                       ("rowband", {"ROWBAND": 8}), ("graded", {"ROWBAND": -2})]:
        for o, v in opts.items():
            mf.set_option(o, v)
        y = _run(mf, xp, k, m, pinned=True)
        assert np.isfinite(y).all(), name
        assert max_rel_err(y, y_ref) <= FP64_TOL, name
        assert mf.last_transfer_bytes() == (span * k * 8, m * k * 8), name
    # resident B (set_b + spmm_resident) takes the same window
    mf.set_b(xp, k); mf.spmm_resident(k)
    y = np.full(m * k, np.nan); mf.fetch_c(y, k)
    assert max_rel_err(y, y_ref) <= FP64_TOL
    # UPLOAD_WINDOW=0: all of B goes up (and the NaNs with it, still unread)
    mf.set_option("UPLOAD_WINDOW", 0); mf.set_option("ROWBAND", 0)
    y = _run(mf, xp, k, m, pinned=True)
    assert max_rel_err(y, y_ref) <= FP64_TOL
    assert mf.last_transfer_bytes() == (c * k * 8, m * k * 8)
    mf.close()


@pytest.mark.parametrize("ngpus", [1, 2, 8])
def test_multi_gpu_own_windows(pkg, L, ngpus):
    """spmm_b200_run_multi on a column-stripe matrix: no shard's window exceeds c / N, so every device uploads its own
    window and runs its own pipeline (no peer gather); a wide matrix keeps the replicated path.  Same y either way."""
    if pkg.load_library().spmm_b200_device_count() < ngpus:
        pytest.skip(f"needs {ngpus} GPUs")
    m, c, k = 6000, 16000, 32
    ia, ja, a = banded_csr(m, 900, 20.0, 80, seed=51)
    ja = (ja + 7000).astype(np.int32)                                      # touched columns: 7000..7899 of 16000
    x = dense_b(c, k, 12)
    xp = x.reshape(k, c).copy(); xp[:, :int(ja.min())] = np.nan; xp[:, int(ja.max()) + 1:] = np.nan; xp = xp.ravel()
    y_ref = L.spmm_oracle(ia, ja, a, m, c, x, k)
    mf = pkg.csr_to_format_multi(ia, ja, a, m, c, len(ja), k, ngpus, variant=pkg.ROWSPLIT)
    buf = np.full(m * k, np.nan); mf.register_host(buf)
    for opts in ({}, {"PIPELINE_MIN_KB": 0}, {"ROWBAND": 8}, {"ROWBAND": -2}, {"UPLOAD_WINDOW": 0, "ROWBAND": -1}):
        for o, v in opts.items():
            mf.set_option(o, v)
        buf[:] = np.nan
        mf.spmm(xp if opts.get("UPLOAD_WINDOW", 1) else x, buf, k)
        assert max_rel_err(buf, y_ref) <= FP64_TOL, (ngpus, opts)
    mf.close()


@pytest.mark.parametrize("variant,k", [("ROWSPLIT", 1), ("ROWSPLIT", 2), ("ROWSPLIT", 4), ("ROWSPLIT", 24), ("ROWSPLIT", 128),
                                       ("MERGE", 1), ("MERGE", 32), ("BCSR", 32), ("BCSR", 128), ("BCSR_FMA", 64)])
def test_no_kernel_reads_b_outside_the_window(pkg, L, variant, k):
    """With the window upload the device copy of B holds the caller's data only between the first and the last touched
    column.  No kernel family may read outside it, not even to multiply by zero (padding slots, dead lanes): x is NaN
    there and the device buffers are poisoned by a first call on an all-NaN operand."""
    m, c = 3000, 8000
    lo, hi = 3100, 3700
    ia, ja, a = banded_csr(m, hi - lo, 14.0, 40, seed=61)
    ja = (ja + lo).astype(np.int32)
    x = dense_b(c, k, 6)
    y_ref = L.spmm_oracle(ia, ja, a, m, c, x, k)
    xp = x.reshape(k, c).copy()
    xp[:, :int(ja.min())] = np.nan; xp[:, int(ja.max()) + 1:] = np.nan
    mf = pkg.csr_to_format(ia, ja, a, m, c, len(ja), k, variant=getattr(pkg, variant))
    mf.set_option("UPLOAD_WINDOW", 0)
    junk = np.full(c * k, np.nan); y = np.empty(m * k)
    mf.spmm(junk, y, k)                                      # d_B / d_x now hold NaN everywhere
    mf.set_option("UPLOAD_WINDOW", 1)
    for opts in ({}, {"PIPELINE_MIN_KB": 0}):
        for o, v in opts.items():
            mf.set_option(o, v)
        y = np.full(m * k, np.nan); mf.spmm(xp.ravel(), y, k)
        assert np.isfinite(y).all(), (variant, k, opts)
        assert max_rel_err(y, y_ref) <= FP64_TOL, (variant, k, opts)
        assert mf.last_transfer_bytes()[0] == (int(ja.max()) - int(ja.min()) + 1) * k * 8
    mf.close()


@pytest.mark.parametrize("variant,k", [("BCSR", 32), ("BCSR", 128), ("BCSR_FMA", 64)])
def test_rowband_blocked_variants(pkg, L, variant, k):
    """The blocked kernels run on row windows too (block-row pointer and C shifted; window starts on multiples of 8 rows):
    same bits as the whole launch, rows not a multiple of 8, empty block rows inside."""
    m, c = 6003, 5000
    ia, ja, a = banded_csr(m, c, 24.0, 150, seed=70 + k, empty_rows=[(1000, 1100)])
    x = dense_b(c, k, 3)
    y_ref = L.spmm_oracle(ia, ja, a, m, c, x, k)
    mf = pkg.csr_to_format(ia, ja, a, m, c, len(ja), k, variant=getattr(pkg, variant))
    assert variant.split("_")[0] in mf.format_name
    mf.set_option("ROWBAND", 0)
    y0 = _run(mf, x, k, m, pinned=False)
    assert max_rel_err(y0, y_ref) <= FP64_TOL
    garbage = dense_b(c, k, 78) * 1e3 - 5e2
    for blocks in (1, 4, 7, 16, 64, -2):
        mf.set_option("ROWBAND", 0)
        _run(mf, garbage, k, m, False)
        mf.set_option("ROWBAND", blocks)
        y = _run(mf, x, k, m, pinned=True)
        assert mf.last_rowband() == (blocks if blocks > 0 else 7), (variant, k, blocks)
        assert np.array_equal(y, y0), (variant, k, blocks)
    mf.close()


@pytest.mark.parametrize("ngpus,variant", [(2, "ROWSPLIT"), (2, "MERGE"), (4, "ROWSPLIT"), (8, "ROWSPLIT"), (8, "BCSR")])
def test_multi_gpu_download_routes(pkg, L, ngpus, variant):
    """MULTI_ROUTE: devices that do not download store their rows of C into a partner's relay buffer over NVLink and the
    partner's copy engine carries them into y.  Forced downloader sets, then the measured choice; same y every time."""
    if pkg.load_library().spmm_b200_device_count() < ngpus:
        pytest.skip(f"needs {ngpus} GPUs")
    m, c, k = 6000, 16000, 32
    ia, ja, a = banded_csr(m, 900, 20.0, 80, seed=52)
    ja = (ja + 7000).astype(np.int32)
    if variant == "MERGE":                        # one long row so that tiles carry into each other inside a relay buffer
        extra = np.setdiff1d(np.arange(7000, 7900, dtype=np.int32), ja[ia[100]:ia[101]])
        row = np.sort(np.concatenate([ja[ia[100]:ia[101]], extra]))
        vals = np.concatenate([a[ia[100]:ia[101]], np.full(len(extra), 0.5)])[np.argsort(np.concatenate([ja[ia[100]:ia[101]], extra]), kind="stable")]
        ja = np.concatenate([ja[:ia[100]], row, ja[ia[101]:]]).astype(np.int32)
        a = np.concatenate([a[:ia[100]], vals, a[ia[101]:]])
        ia = ia.copy(); ia[101:] += len(extra)
    x = dense_b(c, k, 13)
    y_ref = L.spmm_oracle(ia, ja, a, m, c, x, k)
    mf = pkg.csr_to_format_multi(ia, ja, a, m, c, len(ja), k, ngpus, variant=getattr(pkg, variant))
    buf = np.full(m * k, np.nan); mf.register_host(buf)
    masks = [1, 1 << (ngpus - 1), sum(1 << g for g in range(0, ngpus, 2)), sum(1 << g for g in range(ngpus // 2, ngpus)), 0, -1]
    for mask in masks:
        mf.set_option("MULTI_ROUTE", mask)
        buf[:] = np.nan
        mf.spmm(x, buf, k)
        assert max_rel_err(buf, y_ref) <= FP64_TOL, (ngpus, variant, mask)
        route, probe = mf.routes()
        if mask > 0:
            assert all(mask >> r & 1 for r in route), (mask, route)
        elif mask == 0:
            assert route == list(range(ngpus))
        else:
            assert sorted(set(route)) == sorted(r for r in set(route) if route[r] == r)      # every downloader carries itself
            assert (ngpus < 4) == (max(probe) == 0.0)
    mf.close()


def test_routed_downloads_between_shard_handles(pkg, L):
    """spmm_b200_run_routed, the one-process-per-GPU form of download routing, exercised inside one process: three shard
    handles on one device play ranks 0..2, rank 1 downloads for ranks 0 and 2, the transfer callback is a device copy.
    The shared y must equal the oracle's whole result."""
    import torch
    m, c, k = 9000, 12000, 32
    ia, ja, a = banded_csr(m, 1000, 16.0, 90, seed=81)
    ja = (ja + 4000).astype(np.int32)
    x = dense_b(c, k, 14)
    y_ref = L.spmm_oracle(ia, ja, a, m, c, x, k)
    shards = [pkg.csr_to_format(ia, ja, a, m, c, len(ja), k, variant=pkg.ROWSPLIT, rank=r, nranks=3) for r in range(3)]
    y_all = torch.full((m * k,), float("nan"), dtype=torch.float64).pin_memory()
    sent = {}                                   # rank -> tensor: what a client "sent"

    class _Dev:
        def __init__(self, ptr, count):
            self.__cuda_array_interface__ = {"shape": (int(count),), "typestr": "<f8", "data": (int(ptr), False), "version": 3}

    def make_xfer(rank):
        def xfer(ptr, count, peer, is_send, stream):
            with torch.cuda.stream(torch.cuda.ExternalStream(stream)):
                if is_send:
                    sent[rank] = torch.as_tensor(_Dev(ptr, count), device="cuda").clone()     # the message in flight
                else:
                    torch.cuda.current_stream().wait_stream(torch.cuda.default_stream())
                    torch.as_tensor(_Dev(ptr, count), device="cuda").copy_(sent.pop(peer))
        return xfer

    # clients first (their sends complete into `sent`), then the downloader
    for r in (0, 2):
        shards[r].spmm_routed(x, y_all.numpy(), k, r, 1, [], make_xfer(r))
    torch.cuda.synchronize()
    clients = [(r, shards[r].first_row, shards[r].m) for r in (0, 2)]
    shards[1].spmm_routed(x, y_all.numpy(), k, 1, 1, clients, make_xfer(1))
    torch.cuda.synchronize()
    assert max_rel_err(y_all.numpy(), y_ref) <= FP64_TOL
    assert shards[0].last_transfer_bytes()[1] == 0 and shards[1].last_transfer_bytes()[1] == m * k * 8
    # a downloader without clients is the plain call
    y2 = torch.full((m * k,), float("nan"), dtype=torch.float64).pin_memory()
    for r in range(3):
        shards[r].spmm_routed(x, y2.numpy(), k, r, r, [], None)
    assert np.array_equal(y2.numpy(), y_all.numpy())
    for s in shards:
        s.close()
