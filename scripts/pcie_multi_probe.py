#!/usr/bin/env python
"""Aggregate host<->device copy rate of the box when N ranks copy at once (one rank per GPU, torchrun).
Explains the e2e curve of bench.py at N > 1: every rank uploads 1/N of B and downloads its rows of C over its own
PCIe link, but all links end in the same host memory.  Each rank times (CUDA events, after a barrier) `reps` copies of
`mb` MB from / to pinned host memory: H2D alone, D2H alone, both directions at once on two streams.
Rank 0 prints one JSON line per (mb, mode): per-rank GB/s (min over ranks) and the aggregate over the box.
profiles/r2_pcie_multi_probe.md."""
import argparse
import json
import os

import torch
import torch.distributed as dist


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mbs", default="32,128")
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--subsets", default="", help="comma list of rank sets (digits), e.g. 0123,4567,0246: D2H with only those ranks copying")
    args = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    for mb in [int(v) for v in args.mbs.split(",")]:
        n = mb * (1 << 20) // 8
        h_up = torch.empty(n, dtype=torch.float64, pin_memory=True).fill_(1.0)
        h_dn = torch.empty(n, dtype=torch.float64, pin_memory=True)
        d_up = torch.empty(n, dtype=torch.float64, device=dev)
        d_dn = torch.ones(n, dtype=torch.float64, device=dev)
        for mode in ("h2d", "d2h", "both"):
            for timed in (False, True):
                torch.cuda.synchronize()
                if world > 1:
                    dist.barrier()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(s1)
                s2.wait_event(e0)
                for _ in range(args.reps if timed else 3):
                    if mode in ("h2d", "both"):
                        with torch.cuda.stream(s1):
                            d_up.copy_(h_up, non_blocking=True)
                    if mode in ("d2h", "both"):
                        with torch.cuda.stream(s2):
                            h_dn.copy_(d_dn, non_blocking=True)
                s1.wait_stream(s2)
                e1.record(s1)
                torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / args.reps
            gbs = mb * (1 << 20) / (ms * 1e-3) / 1e9 * (2 if mode == "both" else 1)
            t = torch.tensor([gbs], dtype=torch.float64, device=dev)
            lo, tot = t.clone(), t.clone()
            if world > 1:
                dist.all_reduce(lo, op=dist.ReduceOp.MIN)
                dist.all_reduce(tot, op=dist.ReduceOp.SUM)
            if rank == 0:
                print(json.dumps({"ranks": world, "mb_per_copy": mb, "mode": mode, "per_rank_gbs_min": round(float(lo), 1),
                                  "aggregate_gbs": round(float(tot), 1)}), flush=True)
    # which ranks share what: D2H with only a subset of the ranks active (the others wait at the barrier)
    for sub in [v for v in args.subsets.split(",") if v]:
        active = str(rank) in sub
        mb = 32
        n = mb * (1 << 20) // 8
        h_dn = torch.empty(n, dtype=torch.float64, pin_memory=True)
        d_dn = torch.ones(n, dtype=torch.float64, device=dev)
        gbs = 0.0
        for timed in (False, True):
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(s2)
            if active:
                for _ in range(args.reps if timed else 3):
                    with torch.cuda.stream(s2):
                        h_dn.copy_(d_dn, non_blocking=True)
            e1.record(s2)
            torch.cuda.synchronize()
        if active:
            gbs = mb * (1 << 20) / (e0.elapsed_time(e1) / args.reps * 1e-3) / 1e9
        t = torch.tensor([gbs], dtype=torch.float64, device=dev)
        lo = torch.tensor([gbs if active else 1e9], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(lo, op=dist.ReduceOp.MIN)
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
        if rank == 0:
            print(json.dumps({"ranks": world, "active": sub, "mode": "d2h", "mb_per_copy": mb, "per_rank_gbs_min": round(float(lo), 1),
                              "aggregate_gbs": round(float(t), 1)}), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
