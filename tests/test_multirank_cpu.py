"""N > 1 path on CPU: world_size-2 gloo.  The product computes nothing on the CPU, so these tests
cover the host logic of the multi-GPU layout: shard planning with the host twin of
the reference partitioner, B replicated by ONE broadcast, C row-local -- with the ORACLE standing
in for the per-rank kernel (checker only) so that the stitched result can be compared."""
import os
import socket
import sys

import numpy as np
import pytest

from conftest import ROOT, dense_b, random_csr, max_rel_err, FP64_TOL


def test_host_partitioner_matches_oracle_and_pins(pkg, L, pins, golden_cases):
    g = golden_cases[0]
    p = pins[g["args"]]
    ia = g["ia"]; m = len(ia) - 1; nnz = int(ia[-1])
    assert pkg.partition(ia, m, nnz, 8, pkg.PART_NNZ, host=True).tolist() == p["nnz_split_P8"]
    assert pkg.partition(None, m, nnz, 8, pkg.PART_ROWS, host=True).ravel().tolist() == p["rows_split_P8"]
    assert pkg.partition(ia, m, nnz, 8, pkg.PART_MERGE, host=True).ravel().tolist() == p["merge_P8"]
    for seed in range(4):
        mm = 1 + 613 * seed
        ia, ja, a = random_csr(mm, 300, 5.0, seed, empty_frac=0.4, long_row=(seed % 2) * 250)
        for P in (1, 2, 3, 8, 1000):
            o = np.zeros(P + 1, np.int64); L.oracle().oracle_balance_prefix_sums(ia, mm, len(ja), P, o)
            assert np.array_equal(pkg.partition(ia, mm, len(ja), P, pkg.PART_NNZ, host=True), o)
            o = np.zeros(2 * P + 2, np.int64); L.oracle().oracle_merge_partition(ia, mm, len(ja), P, o)
            assert np.array_equal(pkg.partition(ia, mm, len(ja), P, pkg.PART_MERGE, host=True).ravel(), o)


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port, q):
    try:
        sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
        import torch
        import torch.distributed as dist
        import __graft_entry__ as g
        from oracle import loader as L
        pkg = g.load_package()
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
        dist.init_process_group("gloo", rank=rank, world_size=world)
        m, c, k = 3000, 1200, 8
        ia, ja, a = random_csr(m, c, 9.0, seed=5, empty_frac=0.1, long_row=900)   # every rank holds A (as bench.py)
        nnz = len(ja)
        # B exists on rank 0 only; one broadcast replicates it
        x = torch.from_numpy(dense_b(c, k, 3) if rank == 0 else np.zeros(c * k))
        dist.broadcast(x, src=0)
        r0, r1, j0, j1 = pkg.shard_slices(ia, m, nnz, world)[rank]
        sub_ia = (ia[r0:r1 + 1] - ia[r0]).astype(np.int32)
        y_loc = L.spmm_oracle(sub_ia, ja[j0:j1].copy(), a[j0:j1].copy(), r1 - r0, c, x.numpy(), k)   # row-local C
        parts = [None] * world
        dist.all_gather_object(parts, (r0, r1, j1 - j0, y_loc))
        if rank == 0:
            parts.sort(key=lambda t: t[0])
            assert parts[0][0] == 0 and parts[-1][1] == m and all(parts[i][1] == parts[i + 1][0] for i in range(world - 1))
            bounds = np.zeros(world + 1, np.int64); L.oracle().oracle_balance_prefix_sums(ia, m, nnz, world, bounds)
            assert [p[0] for p in parts] + [m] == bounds.tolist()
            assert sum(p[2] for p in parts) == nnz
            y = np.concatenate([p[3] for p in parts])
            y_ref = L.spmm_oracle(ia, ja, a, m, c, dense_b(c, k, 3), k)
            q.put(("ok", max_rel_err(y, y_ref)))
        dist.barrier()
        dist.destroy_process_group()
    except Exception as e:                      # pragma: no cover
        q.put(("err", repr(e)))
        raise


def test_two_rank_gloo_shard_broadcast_stitch(pkg, L):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    status, val = q.get(timeout=240)
    for p in procs:
        p.join(timeout=120)
    assert status == "ok", val
    assert val <= FP64_TOL
    assert all(p.exitcode == 0 for p in procs)
