#!/usr/bin/env python
"""bench.py -- fp64 CSR SpMM (C = A*B) throughput on N B200s vs the reference's CPU kernel.

Metric (BASELINE.json): GFLOP/s = 2*nnz*k / t  (reference formula, spmv_bench.cpp:406) and the
fraction of the HBM roofline (compulsory bytes 12*nnz + 4*(m+1) + 8*k*c + 8*k*m).

  python bench.py [--gpus N --steps K --warmup W]          our arm (CUDA through the C ABI)
  python bench.py --impl reference [...]                    the reference's OpenMP compute_csr
                                                            (oracle/_ref, else the oracle port)
A "step" is one SpMM call over the whole matrix.  Default workload = BASELINE configs[1]
(1M x 1M, avg 50 nnz/row, normal, k=32; generator args in WORKLOADS['cfg2']).  The matrix comes
from the reference's own generator (oracle/_ref/libamg_ref.so, built from /root/reference);
B is seeded uniform [0,1).  Under torchrun rows are sharded nnz-balanced over the ranks, B is
replicated with one NCCL broadcast, C stays row-local: the matrix is fixed, so scaling is strong.
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

HOST_CPUS = len(os.sched_getaffinity(0))     # before libgomp may bind the initial thread to one place

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # generator args in the harness order (spmv_bench.cpp:851-862):
    # rows cols avg std distribution placement bw skew neighbours cross_row_similarity seed
    "cfg1": ("100000 100000 10 2 normal random 1 0 0 0 14", 1),
    "cfg2": ("1000000 1000000 50 16.6667 normal random 0.05 0 0.05 0.05 14", 32),
    "cfg2_bw0.3": ("1000000 1000000 50 16.6667 normal random 0.3 0 0.05 0.05 14", 32),
    "cfg2_bw1": ("1000000 1000000 50 16.6667 normal random 1 0 0.05 0.05 14", 32),
    "cfg3a": ("2000000 2000000 20 20 gamma random 0.3 10000 0.5 0.5 14", 32),
    "cfg3b": ("2000000 2000000 20 40 gamma random 0.3 0 0.5 0.5 14", 32),
    "cfg4": ("1000000 1000000 50 16.6667 normal diagonal 0.0002 0 1.9 0.95 14", 128),
    "cfg5_large": ("8388607 8388607 5 1.6667 normal random 0.05 0 0.05 0.05 14", 128),
    "small": ("50000 50000 20 6 normal random 0.3 0 0.05 0.05 14", 32),
}
METRIC = "spmm_fp64_csr_gflops"
UNIT = "GFLOP/s"


def algorithmic_bytes(nnz, m, c, k):
    return 12.0 * nnz + 4.0 * (m + 1) + 8.0 * k * c + 8.0 * k * m


def dense_b(c, k, seed=42):
    """Seeded uniform [0,1) B in the oracle layout x[n*c + j]."""
    return np.random.default_rng(seed).random(c * k)


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic(workload, k, world, kernel):
    """dram bytes per launch of the dominant kernel from the committed ncu capture (profiles/traffic.json), or None.
    Only for the configuration the capture was taken on: one GPU, same workload and k, same kernel family."""
    if world != 1:
        return None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f)
        e = t.get(f"{workload}_k{k}")
        if isinstance(e, dict):
            return e.get("dram_bytes") if e.get("kernel", kernel) == kernel else None
        return e
    except Exception:
        return None


def cpu_model():
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("model name"):
                    return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


# L2 -> SM gather ceiling measured with scripts/gather_probe.cu on this pool's B200 (profiles/r1_gather_probe.log):
# random 256-byte rows out of an L2-resident window
L2_GATHER_CEILING_TBS = 21.6


class ClockSampler:
    """Samples SM clocks and throttle reasons with NVML while the timed region runs."""

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._t = None
        try:
            import pynvml as N
            N.nvmlInit()
            self.N = N
            self.h = N.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = N.nvmlDeviceGetMaxClockInfo(self.h, N.NVML_CLOCK_SM)
        except Exception:
            self.N = None

    def _loop(self):
        N = self.N
        names = {}
        for nm in dir(N):
            if nm.startswith("nvmlClocksThrottleReason") or nm.startswith("nvmlClocksEventReason"):
                v = getattr(N, nm)
                if isinstance(v, int) and v not in (0,) and bin(v).count("1") == 1:
                    names.setdefault(v, nm.replace("nvmlClocksThrottleReason", "").replace("nvmlClocksEventReason", ""))
        while not self._stop.is_set():
            try:
                self.samples.append(N.nvmlDeviceGetClockInfo(self.h, N.NVML_CLOCK_SM))
                try:
                    r = N.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = N.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, nm in names.items():
                    if r & bit and "Idle" not in nm and nm not in ("None", "All"):
                        self.reasons.add(nm)
            except Exception:
                pass
            time.sleep(0.01)

    def start(self):
        if self.N:
            self._t = threading.Thread(target=self._loop, daemon=True)
            self._t.start()

    def stop(self):
        self._stop.set()
        if self._t:
            self._t.join()
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


# ---------------------------------------------------------------------------------------------
# CPU legs (the only places this file executes oracle/)
# ---------------------------------------------------------------------------------------------

def use_all_host_threads():
    """torchrun exports OMP_NUM_THREADS=1 to every rank; the CPU arm must use the cores the process may run on."""
    import ctypes
    n = HOST_CPUS
    try:
        ctypes.CDLL("libgomp.so.1").omp_set_num_threads(n)
    except OSError:
        pass
    return n


def cpu_spmm_timer(ia, ja, a, m, c, k, x):
    """Returns (fn(rows)->seconds, kind, cores): one call of the reference CPU kernel on the first
    `rows` rows.  kind = 'reference' when oracle/_ref holds the compiled reference."""
    from oracle import loader as L
    if L.have_ref():
        L.ref()
    else:
        L.oracle()
    use_all_host_threads()          # after the library (and its libgomp) is loaded
    y = np.empty(m * k)
    if L.have_ref():
        lib, kind, cores = L.ref(), "reference", L.ref().ref_num_threads()

        def run(rows, first=0):
            # rows [first, first + rows): row_ptr holds absolute offsets into ja / a, so a row window is ia + first
            nz = int(ia[first + rows])
            t0 = time.perf_counter()
            lib.ref_spmm_csr_nocopy(ia[first:], ja, a, rows, c, nz, x, y, k)
            return time.perf_counter() - t0
    else:
        lib, kind, cores = L.oracle(), "port", L.oracle().oracle_num_threads()

        def run(rows, first=0):
            t0 = time.perf_counter()
            lib.oracle_spmm_csr(ia[first:], ja, a, rows, c, x, y, k)
            return time.perf_counter() - t0
    run.y = y
    return run, kind, cores


def bounded_rows(run, ia, m, k, budget_s):
    """Pick a row prefix so that one call costs about budget_s (full matrix if it is cheaper)."""
    probe = max(1, min(m, 20000))
    run(probe)
    t = max(run(probe), 1e-6)
    rate = 2.0 * float(ia[probe]) * k / t          # flop/s on the probe
    full = 2.0 * float(ia[m]) * k / rate
    if full <= budget_s:
        return m
    target_nnz = rate * budget_s / (2.0 * k)
    return int(max(probe, min(m, np.searchsorted(ia, target_nnz))))


def reference_parity(ia, ja, a, c, k, x, y_gpu, first_row, rows):
    """max_i |y - y_ref| / max(|y_ref|, tiny) of the GPU result (host buffer the e2e call filled) against the
    reference's own compute_csr (spmm_kernel_csr.cpp:69-97) on the same rows of the same matrix and the same B."""
    run, kind, _ = cpu_spmm_timer(ia, ja, a, len(ia) - 1, c, k, x)
    run(rows, first_row)
    y_ref = run.y[: rows * k]
    if rows == 0:
        err = 0.0
    else:
        err = float(np.max(np.abs(y_gpu[: rows * k] - y_ref) / np.maximum(np.abs(y_ref), 1e-300)))
    return {"max_rel_err": err, "tol": 1e-12, "ok": bool(err <= 1e-12), "rows_compared": int(rows), "first_row": int(first_row),
            "vs": "reference compute_csr (oracle/_ref/libspmm_ref.so)" if kind == "reference" else "oracle port (oracle/liboracle.so)",
            "against": "y returned by the e2e host-buffer call"}


def cpu_baseline(ia, ja, a, m, c, k, x, total_budget_s=20.0):
    run, kind, cores = cpu_spmm_timer(ia, ja, a, m, c, k, x)
    rows = bounded_rows(run, ia, m, k, budget_s=total_budget_s / 4.0)
    nz = float(ia[rows])
    times, t_start = [], time.perf_counter()
    run(rows)                                        # warm-up
    while len(times) < 3 or (time.perf_counter() - t_start < total_budget_s and len(times) < 20):
        times.append(run(rows))
    t = float(np.median(times))
    sample = (f"full matrix ({rows} rows, {int(nz)} nnz)" if rows == m else
              f"first {rows} of {m} rows ({int(nz)} nnz)") + f", {len(times)} timed calls after 1 warm-up, median"
    return {"value": 2.0 * nz * k / t / 1e9, "unit": UNIT, "cores": int(cores), "kind": kind, "sample": sample,
            "cpu_model": cpu_model(), "omp": {v: os.environ.get(v) for v in ("OMP_PROC_BIND", "OMP_PLACES", "OMP_DYNAMIC")}}


def cpu_baseline_subprocess(args, k):
    """The cpu_baseline leg as its own process = `bench.py --impl reference` on the same workload: the reference pins
    one OpenMP thread per core (OMP_PROC_BIND / OMP_PLACES, run.sh:618-619), which must be in the environment before
    libgomp starts and would also pin this process's own host thread."""
    import subprocess
    cmd = [sys.executable, os.path.abspath(__file__), "--impl", "reference", "--workload", args.workload, "--k", str(k),
           "--steps", "5", "--warmup", "1"]
    env = {kk: v for kk, v in os.environ.items() if kk not in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "OMP_NUM_THREADS")}
    try:
        r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)
        for ln in reversed(r.stdout.strip().splitlines()):
            if ln.startswith("{"):
                return json.loads(ln)["cpu_baseline"]
    except Exception:
        pass
    return None


def reference_arm(args, ia, ja, a, m, c, k, nnz, workload_desc):
    """--impl reference: the reference's own CPU implementation, all host threads."""
    x = dense_b(c, k)
    run, kind, cores = cpu_spmm_timer(ia, ja, a, m, c, k, x)
    steps_total = args.steps + args.warmup
    rows = bounded_rows(run, ia, m, k, budget_s=max(0.05, 150.0 / max(steps_total, 1)))
    nz = float(ia[rows])
    for _ in range(args.warmup):
        run(rows)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        run(rows)
    t = (time.perf_counter() - t0) / max(args.steps, 1)
    val = 2.0 * nz * k / t / 1e9
    sample = (f"full matrix per step ({int(nz)} nnz)" if rows == m else
              f"first {rows} of {m} rows per step ({int(nz)} of {nnz} nnz)")
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": t * 1e3, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_desc,
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": int(cores), "kind": kind, "sample": sample,
                             "cpu_model": cpu_model(),
                             "omp": {v: os.environ.get(v) for v in ("OMP_PROC_BIND", "OMP_PLACES", "OMP_DYNAMIC")}},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------------------------

def measure_download_routes(torch, dist, dev, rank, world):
    """One process per GPU: which ranks should carry C to the host on THIS box.  Times device->host copies (16 MB, copy
    engine) with all ranks active and with four candidate subsets; when a subset moves more than 1.2x what all ranks
    move together, the other ranks route their rows of C through a partner in it.  Returns (route, probe):
    route[r] = downloader of rank r (identity when nothing is gained), probe = GB/s of {all, lower, upper, even, odd}."""
    ident = list(range(world))
    if world < 4:
        return ident, None
    n = (16 << 20) // 8
    h = torch.empty(n, dtype=torch.float64, pin_memory=True)
    d = torch.ones(n, dtype=torch.float64, device=dev)
    sets = [ident, ident[: world // 2], ident[world // 2:], ident[0::2], ident[1::2]]
    rates = []
    for si, members in enumerate([ident] + sets):          # the first pass is a warm-up
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
        t0 = time.perf_counter()
        if rank in members:
            for _ in range(3):
                h.copy_(d, non_blocking=True)
            torch.cuda.synchronize()
        el = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        dist.all_reduce(el, op=dist.ReduceOp.MAX)
        if si > 0:
            rates.append(3 * len(members) * n * 8 / float(el.item()) / 1e9)
    best = max(range(len(sets)), key=lambda i: rates[i])
    if best == 0 or rates[best] <= 1.2 * rates[0]:
        return ident, rates
    down = sets[best]
    route, nxt = [], 0
    for r in ident:
        if r in down:
            route.append(r)
        else:
            route.append(down[nxt % len(down)]); nxt += 1
    return route, rates


def gpu_arm(args, ia, ja, a, m, c, k, nnz, workload_desc):
    import torch
    import torch.distributed as dist
    import __graft_entry__ as g
    pkg = g.load_package()

    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product has no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    variant = {"auto": pkg.AUTO, "rowsplit": pkg.ROWSPLIT, "merge": pkg.MERGE, "bcsr": pkg.BCSR,
               "bcsr_fma": pkg.BCSR_FMA}[args.variant]
    if world > 1:
        mf = pkg.csr_to_format(ia, ja, a, m, c, nnz, k, device=local, variant=variant, rank=rank, nranks=world)
    else:
        mf = pkg.csr_to_format(ia, ja, a, m, c, nnz, k, device=local, variant=variant)
    m_loc = mf.m

    # B: every rank holds the seeded host copy (pinned, oracle layout).  Device-resident leg: rank 0 uploads, one NCCL
    # broadcast replicates.  e2e leg: every rank uploads its 1/N part over its own PCIe link, the parts are exchanged
    # over NVLink (one broadcast per owner = all-gather)
    x_host = torch.empty(c * k, dtype=torch.float64, pin_memory=True)
    x_host.numpy()[:] = dense_b(c, k)        # SPMD: every rank generates the (seeded) inputs, like A above
    y_host = torch.empty(max(m_loc * k, 1), dtype=torch.float64, pin_memory=True)
    x_dev = torch.empty(c * k, dtype=torch.float64, device=dev)      # column-major staging
    b_dev = torch.empty(c * k, dtype=torch.float64, device=dev)      # row-major operand
    c_dev = torch.empty(max(m_loc * k, 1), dtype=torch.float64, device=dev)
    stream = torch.cuda.current_stream()

    def replicate_b():
        if rank == 0:
            x_dev.copy_(x_host, non_blocking=True)
        if world > 1:
            dist.broadcast(x_dev, src=0)
        if k == 1:
            b_dev.copy_(x_dev)
        else:
            mf.transpose_b_device(x_dev.data_ptr(), b_dev.data_ptr(), k, stream.cuda_stream)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v):
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    replicate_b()
    barrier()

    # ---- device-resident timing: `value` -----------------------------------------------------
    for _ in range(max(args.warmup, 9) if k >= 32 else args.warmup):      # k >= 32: room for the k-slice measurement calls (large windows only)
        mf.spmm_device(b_dev.data_ptr(), c_dev.data_ptr(), k, stream.cuda_stream)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    sampler = ClockSampler(local)
    launches0 = mf.launch_count()
    barrier()
    sampler.start()
    ev[0].record(stream)
    for i in range(args.steps):
        mf.spmm_device(b_dev.data_ptr(), c_dev.data_ptr(), k, stream.cuda_stream)
        ev[i + 1].record(stream)
    barrier()
    launches = mf.launch_count() - launches0
    total_ms = ev[0].elapsed_time(ev[-1])
    per_step = [ev[i].elapsed_time(ev[i + 1]) for i in range(args.steps)]
    total_ms = max_over_ranks(total_ms)
    kernel_ms = max_over_ranks(float(np.mean(per_step)))     # back-to-back launches: step = kernel(s)
    ms_per_step = total_ms / args.steps
    value = 2.0 * nnz * k / (ms_per_step * 1e-3) / 1e9

    # ---- end to end with HOST buffers: `e2e` -----------------------------------------------------
    e2e_steps = max(3, min(args.steps, 10))

    class _DevArray:           # library-owned device range -> torch tensor (for the broadcast callback)
        def __init__(self, ptr, n):
            self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f8", "data": (ptr, False), "version": 2}

    def bcast(ptr, count, parts, cuda_stream):
        t = torch.as_tensor(_DevArray(ptr, count), device=dev)
        with torch.cuda.stream(torch.cuda.ExternalStream(cuda_stream, device=dev)):
            if parts == 1:
                dist.broadcast(t, src=0)
            elif count % parts == 0:          # every rank's part has the same length: one all-gather = P broadcasts
                per = count // parts
                dist.all_gather_into_tensor(t, t[rank * per:(rank + 1) * per])
            else:
                per = (count + parts - 1) // parts
                works = [dist.broadcast(t[r * per:min(count, (r + 1) * per)], src=r, async_op=True)
                         for r in range(parts) if r * per < count]
                for w in works:
                    w.wait()

    def xfer(ptr, count, peer, is_send, lib_stream):
        # send / receive of C rows between ranks for routed downloads: NCCL point-to-point on the library's stream
        with torch.cuda.stream(torch.cuda.ExternalStream(lib_stream, device=dev)):
            t = torch.as_tensor(_DevArray(ptr, count), device=dev)
            if is_send:
                dist.send(t, dst=peer)
            else:
                dist.recv(t, src=peer)

    def e2e_step():
        # the public drop-in call with HOST buffers: upload (root) -> broadcast -> kernel -> download of the
        # row-local C, pipelined inside the library over row blocks (banded matrix, N = 1) or 16-column slices
        if routed:
            mf.spmm_routed(x_host.numpy(), y_all.numpy(), k, rank, route[rank], clients, xfer)
        elif world == 1 or own_window:
            mf.spmm(x_host.numpy(), y_host.numpy()[: m_loc * k], k)
        else:
            mf.spmm_replicated(x_host.numpy(), y_host.numpy()[: m_loc * k], k, bcast, x_on_every_rank=True)

    def time_e2e(steps):
        for _ in range(2):
            e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            e2e_step()
        barrier()
        return max_over_ranks((time.perf_counter() - t0) / steps)

    # Every rank holds x (the reference's harness is one process with one x).  A shard reads only the rows of B between
    # its first and last touched column; when that window is about the 1/N share of B a rank would upload for
    # the all-gather, every rank uploads its own window over its own PCIe link and no collective runs at all.
    span = mf.col_window()[1]
    own_window = world > 1 and max_over_ranks(float(span)) * world <= 1.25 * c      # same rule as spmm_b200_run_multi
    # Download routing: the PCIe links of a box are not equal.  With own windows, N >= 4 and a result small
    # enough for one shared-memory y, the ranks measure which of them should carry C to the host; the others send their
    # rows to a partner over NVLink (NCCL send / recv) and the partner stores them into the shared y.
    route, route_probe, routed, clients, y_all, shm_path = list(range(world)), None, False, [], None, None
    shm_free = 0
    if world > 1 and os.path.isdir("/dev/shm"):
        st = os.statvfs("/dev/shm")
        shm_free = -int(max_over_ranks(-float(st.f_bavail * st.f_frsize)))          # min over ranks
    if own_window and world >= 4 and 8 * m * k + (64 << 20) <= min(2 << 30, shm_free) and not os.environ.get("SPMM_B200_BENCH_NO_ROUTE"):
        route, route_probe = measure_download_routes(torch, dist, dev, rank, world)
        routed = any(route[r] != r for r in range(world))
    forced = os.environ.get("SPMM_B200_BENCH_FORCE_ROUTE")          # e.g. "1,1": rank 1 downloads for both (testing at N = 2)
    if forced and own_window and world > 1:
        route = [int(v) for v in forced.split(",")]
        assert len(route) == world and all(route[route[r]] == route[r] for r in range(world)), "a downloader must carry its own rows"
        routed = any(route[r] != r for r in range(world))
    if routed:
        first = torch.tensor([mf.first_row if r == rank else 0 for r in range(world)] + [m_loc if r == rank else 0 for r in range(world)],
                             dtype=torch.int64, device=dev)
        dist.all_reduce(first, op=dist.ReduceOp.SUM)
        first = first.tolist()
        clients = [(r, first[r], first[world + r]) for r in range(world) if route[r] == rank and r != rank]
        shm_path = "/dev/shm/spmm_b200_bench_y_%s" % os.environ.get("MASTER_PORT", "0")
        if rank == 0:
            with open(shm_path, "wb") as f:
                f.truncate(8 * m * k)
        dist.barrier()
        y_all = torch.from_file(shm_path, shared=True, size=m * k, dtype=torch.float64)      # ONE y for all ranks
        mf.register_host(y_all.numpy())
        dist.barrier()
        if rank == 0:
            os.unlink(shm_path)              # the mappings keep it alive
    e2e_s = time_e2e(e2e_steps)
    if routed:
        y_host = y_all[mf.first_row * k: (mf.first_row + m_loc) * k]       # this rank's rows of the shared result
    clocks = sampler.stop()
    e2e_val = 2.0 * nnz * k / e2e_s / 1e9
    timeline = mf.last_timeline()            # event times of the last e2e step on this rank (ms since its first enqueue)
    band = mf.last_rowband()
    up_bytes, down_bytes = mf.last_transfer_bytes()       # what the library actually moved over PCIe on this rank
    if world > 1:
        t = torch.tensor([up_bytes, down_bytes], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        up_bytes, down_bytes = int(t[0].item()), int(t[1].item())
    # the same call with the whole of B uploaded and the k-slice pipeline (what round 1 and the first half of round 2
    # measured): kept on the line so that the effect of uploading only the touched rows of B is visible
    full = None
    if world == 1 and up_bytes < 8 * c * k:
        y_keep = y_host.clone()
        mf.set_option("UPLOAD_WINDOW", 0); mf.set_option("ROWBAND", 0)
        full_s = time_e2e(3)
        full = {"ms_per_step": full_s * 1e3, "value": 2.0 * nnz * k / full_s / 1e9, "h2d_bytes_per_step": mf.last_transfer_bytes()[0],
                "same_result": bool(torch.equal(y_keep, y_host)),
                "note": "UPLOAD_WINDOW=0, ROWBAND=0: all c x k of B uploaded, k-slice pipeline"}
        mf.set_option("UPLOAD_WINDOW", 1); mf.set_option("ROWBAND", -1)
        y_host.copy_(y_keep)

    # ---- checksum so that the work cannot be skipped silently -----------------------------------
    csum = float(c_dev[: m_loc * k].sum().item()) if m_loc else 0.0
    csum_e2e = float(y_host[: m_loc * k].sum().item()) if m_loc else 0.0      # what the host-buffer call returned
    if abs(csum_e2e - csum) > 1e-9 * max(abs(csum), 1.0):
        raise SystemExit(f"bench.py: e2e result differs from the device-resident result ({csum_e2e} vs {csum})")

    peak, peak_src = measured_peak_gbs()
    nnz_loc = mf.nnz
    # Bytes one launch must move.  N = 1: SURVEY 8(d)'s formula (all of B charged once).  N > 1: a shard only reads the
    # B rows its columns touch (random placement: the same window on every rank), so the per-rank figure charges the
    # touched window (first..last touched column) instead of all of B -- otherwise the "fraction" exceeds 1.
    j0, j1 = int(ia[mf.first_row]), int(ia[mf.first_row + m_loc])
    if j1 > j0:
        col_lo, col_hi = int(ja[j0:j1].min()), int(ja[j0:j1].max())
        touched_b = 8.0 * k * (col_hi - col_lo + 1)
    else:
        touched_b = 0.0
    bytes_full = algorithmic_bytes(nnz_loc, m_loc, c, k)
    bytes_launch = bytes_full if world == 1 else 12.0 * nnz_loc + 4.0 * (m_loc + 1) + touched_b + 8.0 * k * m_loc
    achieved = bytes_launch / (kernel_ms * 1e-3) / 1e9
    gather_bytes = 8.0 * k * nnz_loc         # one B row per nonzero crosses L2 -> SM
    gather_tbs = gather_bytes / (kernel_ms * 1e-3) / 1e12
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic", "config": workload_desc,
        "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": up_bytes, "d2h_bytes_per_step": down_bytes,
                "ms_per_step": e2e_s * 1e3, "steps": e2e_steps,
                "x_bytes": 8 * c * k, "b_rows_uploaded": "first..last column the matrix touches (%d of %d rows of B on rank 0); "
                                                          "rows no nonzero references cannot change y" % (span, c)},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": ncu_traffic(args.workload, k, world, mf.format_name), "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": bytes_launch, "kernel_ms": kernel_ms,
                     "kernel": mf.format_name, "frac_of_8TBs_nominal": achieved / 8000.0,
                     "touched_b_bytes": touched_b,
                     "bytes_model": "12*nnz + 4*(m+1) + 8*k*c + 8*k*m (SURVEY 8(d))" if world == 1 else
                                    "rank 0 shard: 12*nnz_g + 4*(m_g+1) + touched B window + 8*k*m_g",
                     "l2_gather": None if "BCSR" in mf.format_name else {"bytes": gather_bytes, "achieved_tbs": gather_tbs, "ceiling_tbs": L2_GATHER_CEILING_TBS,
                                   "frac": gather_tbs / L2_GATHER_CEILING_TBS,
                                   "note": "nnz*8k bytes of B rows cross L2->SM per launch; ceiling = pure gather of random "
                                           "256-byte rows from an L2-resident window (scripts/gather_probe.cu); it is the "
                                           "binding roof when rows share no columns"},
                     "rank0_shard": world > 1},
        "clocks": clocks,
        "checksum_c_rank0": csum, "checksum_c_e2e_rank0": csum_e2e,
        "variant": mf.format_name,
        "bcsr_fill": mf.bcsr_fill(), "kslice": mf.kslice(),
    }
    line["e2e"]["timeline_ms_last_step_rank0"] = timeline
    # which host pipeline the library took for the e2e call: row blocks of a banded matrix (B rows up, C rows down)
    # or column slices
    line["e2e"]["pipeline"] = (f"row-band, {band} row blocks" if band > 0 else
                               f"k-slice, {len(timeline)} column slices" if timeline else "single upload / kernel / download")
    if world > 1:
        line["e2e"]["replication"] = ("none: every rank uploads its own window of B" if own_window else
                                      "1/N of B per rank + NCCL all-gather per slice")
        line["e2e"]["download_routes"] = {"route": route, "probe_gbs_all_lower_upper_even_odd": route_probe,
                                          "note": "route[r] = rank whose PCIe link carries rank r's rows of C into the shared y "
                                                  "(identity: every rank its own rows); measured at set-up"}
        if routed:
            line["e2e"]["pipeline"] = "routed: own window up, shard kernel, C rows to the downloader rank over NVLink, one contiguous copy per shard"
    if full:
        line["e2e"]["full_b_upload"] = full
    if rank == 0 and not args.no_parity:
        # the e2e result (host buffer) against the reference's CPU kernel on the same rows: whole matrix at N = 1,
        # rank 0's shard rows at N > 1
        line["parity"] = reference_parity(ia, ja, a, c, k, x_host.numpy(), y_host.numpy(), mf.first_row, m_loc)
    if rank == 0 and world == 1 and not args.no_cpu:
        line["cpu_baseline"] = cpu_baseline_subprocess(args, k) or cpu_baseline(ia, ja, a, m, c, k, x_host.numpy())
    if rank == 0:
        print(json.dumps(line), flush=True)
        if "parity" in line and not line["parity"]["ok"]:
            raise SystemExit(f"bench.py: GPU result differs from the reference: {line['parity']}")
    mf.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="cfg2", choices=sorted(WORKLOADS))
    ap.add_argument("--k", type=int, default=None)
    ap.add_argument("--variant", default="auto", choices=["auto", "rowsplit", "merge", "bcsr", "bcsr_fma"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-parity", action="store_true", help="skip the comparison of the e2e result with the reference CPU kernel")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    rank = int(os.environ.get("RANK", 0))
    if args.impl == "reference" and rank != 0:
        return 0
    if args.impl == "reference":
        # the reference pins one OpenMP thread per core (run.sh:618-619); must be set before libgomp starts
        os.environ.setdefault("OMP_PROC_BIND", "true")
        os.environ.setdefault("OMP_PLACES", "cores")
        os.environ.setdefault("OMP_DYNAMIC", "false")

    from oracle import loader as L          # input generation = the reference's own generator
    gen_args, k = WORKLOADS[args.workload]
    if args.k:
        k = args.k
    if not L.have_ref():
        raise SystemExit("bench.py: oracle/_ref/libamg_ref.so (the reference's matrix generator) is missing; "
                         "run __graft_entry__.build() in the container that has /root/reference")
    L.amg()
    use_all_host_threads()              # the generator's output does not depend on the thread count
    ia, ja, a, m, c = L.generate(gen_args)
    nnz = len(ja)
    desc = {"workload": f"{args.workload}: artificial-matrix-generator '{gen_args}', fp64 CSR SpMM k={k}",
            "m": m, "c": c, "nnz": nnz, "k": k, "b_layout": "oracle column-major x[n*c+j] at the boundary, row-major on device",
            "l2": "inputs larger than L2 (A stream + C = %.0f MB per step >> 126 MB); no flush" % ((12.0 * nnz + 8.0 * k * m) / 1e6),
            "sharding": "nnz-balanced rows (lib/parallel_util.h:141-165), B replicated on device (e2e: see e2e.replication), C row-local"}
    if args.impl == "reference":
        reference_arm(args, ia, ja, a, m, c, k, nnz, desc)
    else:
        gpu_arm(args, ia, ja, a, m, c, k, nnz, desc)
    return 0


if __name__ == "__main__":
    sys.exit(main())
