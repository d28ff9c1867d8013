#!/usr/bin/env python
"""bench.py -- LQR solves/sec for the batched quadrotor configuration (BASELINE.json configs[2]:
n=12, m=4, N=256, batch 4096 per GPU, synthetic generator of SURVEY App. D, seed 2024).

One "step" = one pass of the hot path (pack rhs + leaves + K factor levels + rhs propagation) over
the whole batch. `value` is measured with the problems already resident in HBM; `e2e` goes through
the C-ABI with HOST buffers (pinned), host->device copy of every problem and device->host copy of
every solution inside the timed region.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--batch B] [--mode auto|staged|fused]

Under torchrun (N > 1) every rank owns one GPU and its own shard of `batch` problems (weak scaling,
no data-path collective); timing is bracketed by barriers and reduced with MAX over ranks.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
N_STATES, N_INPUTS, HORIZON, SEED = 12, 4, 256, 2024
METRIC = "LQR solves/sec (n=12,m=4,N=256 batched)"
VALUE_KIND = ("resident: problems already in HBM when the timed region starts, solutions left in HBM; "
              "the host-to-host figure through the C-ABI is `e2e` (the one to hold against the reference arm)")
WORKLOAD = "batched quadrotor LQR n=12 m=4 N=256 (BASELINE configs[2])"


# ------------------------------------------------------------------------------------------
def flop_model(n, m, N):
    """Structurally non-zero FLOPs of one solve."""
    K = N.bit_length() - 1
    t_ch = N - 1
    t_cs = 2 * N - 2 * K - 2
    t_ip = t_cs + (N - 1)
    t_sc = 2 * N * (K - 1) - (2 ** (K + 1) - 4)
    t_scl = 0
    for j in range(K - 1):
        size = 2 ** (j + 1)
        for leaf in range(N // size):
            a, b = leaf * size, leaf * size + size - 1
            cols = (a > 0) + (b < N - 1)
            t_scl += cols * (size - 2 + (1 if a == 0 else 0))
    leaf = N * (4 * n ** 3 + n ** 3 / 3 + m ** 3 / 3 + 2 * m * m * n)
    return (t_ip * (4 * n ** 3 + 4 * m * n * n) + t_ch * n ** 3 / 3 + t_cs * 2 * n ** 3 +
            t_sc * (2 * n ** 3 + 2 * m * n * n) + t_scl * 2 * n ** 3 + leaf)


def tier0_flops_per_group(n, m, G):
    """FLOPs tier0_kernel executes for one group of G knots: dense Schur update of every
    knot at each of the log2(G) levels (rows 2n+m, columns [L|R|z] = 2n+1, inner dimension n) plus per separator
    the inner products [A|B](n x (n+m)) x (n+m x 2n+1), the LDL^T factorization (n^3/6) and 2n+1 solves (n^2 each)."""
    g = G.bit_length() - 1
    zs, w = 2 * n + m, 2 * n + 1
    schur = g * G * zs * w * n
    spine = (G - 1) * (n * (n + m) * w + n ** 3 / 6 + w * n * n)
    return 2.0 * (schur + spine)


def scan_leaf_flops_per_subtree(n, m):
    """FLOPs scan_leaf_kernel executes for one 4-knot subtree: two level-0 merges formed from the
    flat knot blocks (S = [A B] D [A B]^T, s, S_L; LDL^T; 2n+1 solves; A^T [f_L f_R zs]; row scalings) and one general
    merge (inner products [A|B] x (n+m x 2n+1), LDL^T, 2n+1 solves, lean Schur update of the 2n+m boundary rows)."""
    w, zs = 2 * n + 1, 2 * n + m
    raw = n * n * (n + m) + n * (n + m) + n * n + n ** 3 / 6 + w * n * n + n * n * w + (n + m) * w
    gen = n * (n + m) * w + n ** 3 / 6 + w * n * n + zs * w * n
    return 2.0 * (2 * raw + gen)


def scan_leaf_bytes_per_subtree(n, m, dual=False):
    """Algorithmic HBM bytes of the leaf kernel per 4-knot subtree: 4 flat knot blocks read; one entry and three
    separator records [f_L | f_R | zs] written. The entry is the (2n+m)-row boundary entry of kernels_scan.cuh, or
    (dual) the Schur-complement element of kernels_dual.cuh: 3 blocks of np x (np+2) + 2 np, np = n rounded up to 4."""
    stride = 2 * n + 2 * m + 1 + n * n + n * m + n
    np_ = (n + 3) & ~3
    ent = 3 * np_ * (np_ + 2) + 2 * np_ if dual else (2 * (2 * n + m) * n + (2 * n + m) + n + 1) & ~1
    return 8 * (4 * stride + ent + 3 * (2 * n + 1) * n)


def whole_step_bytes(n, m, N, dual):
    """Algorithmic HBM bytes of one whole scan-path solve: the leaf kernel's bytes, every tier's elements in and out and its
    records, the top-down pass (records + multipliers) and the down pass (records, knot blocks, multipliers in, solution out)."""
    stride = 2 * n + 2 * m + 1 + n * n + n * m + n
    np_ = (n + 3) & ~3
    ent = 3 * np_ * (np_ + 2) + 2 * np_ if dual else (2 * (2 * n + m) * n + (2 * n + m) + n + 1) & ~1
    wsz = (2 * n + 1) * n
    M0 = N // 4
    leaf = M0 * (4 * stride + ent + 3 * wsz)
    tiers, M = 0, M0
    while M > 1:
        e = min(4, M)
        tiers += M * ent + (M // e) * ent + (M - M // e) * wsz
        M //= e
    top = (M0 - 1) * wsz + 2 * M0 * n
    down = M0 * (3 * wsz + 4 * stride + 2 * n) + N * (2 * n + m)
    return 8 * (leaf + tiers + top + down)


# dram__bytes_read.sum + dram__bytes_write.sum of scan_leaf_kernel per problem from the committed ncu --set full capture
# (profiles/r02_ncu_full.md: 497.2 MB + 698.1 MB at batch 1024, hybrid leaf kernel)
SCAN_LEAF_TRAFFIC_BYTES_PER_PROBLEM = (497.2e6 + 698.1e6) / 1024

# dram__bytes_read.sum + dram__bytes_write.sum of tier0_kernel per problem, from the committed ncu --set full capture
# (profiles/r01_ncu_tier0_v1.md: 0.995 GB + 4.060 GB at batch 2048)
TIER0_TRAFFIC_BYTES_PER_PROBLEM = (0.995207936e9 + 4.060259e9) / 2048


def fp64_peak():
    """fp64 tensor (DMMA m8n8k4) peak measured on this pool's B200 by tools/fp64_peak.cu (MEASURED_PEAKS.json has
    no fp64 entry)."""
    p = os.path.join(ROOT, "profiles", "r01_fp64_peak.json")
    try:
        return {"tflops": float(json.load(open(p))["dmma_m8n8k4_tflops"]), "kind": "measured (tools/fp64_peak.cu, profiles/r01_fp64_peak.json)"}
    except Exception:
        return {"tflops": 37.2, "kind": "fallback: 148 SM x 64 FMA/clk x 1.965 GHz"}


def schur_bytes_model(n, m, N, batch):
    """Algorithmic HBM bytes of the staged Schur launches, averaged per launch: every active task
    (knot, boundary column) reads F(k,level) and reads+writes F(k,column) (SURVEY 8d streaming model)."""
    K = N.bit_length() - 1
    fs = (2 * n + m) * n
    tot = 0
    for j in range(K - 1):
        size = 2 ** (j + 1)
        for leaf in range(N // size):
            a, b = leaf * size, leaf * size + size - 1
            cols = (a > 0) + (b < N - 1)
            tot += size * (fs * (1 if cols else 0) + 2 * fs * cols) * 8
    return tot * batch / (K - 1)


class stdout_to_stderr:
    """NCCL prints its version banner on fd 1 at the first collective; keep stdout = the one JSON line."""

    def __enter__(self):
        sys.stdout.flush()
        self.saved = os.dup(1)
        os.dup2(2, 1)

    def __exit__(self, *exc):
        sys.stdout.flush()
        os.dup2(self.saved, 1)
        os.close(self.saved)


def bind_to_gpu_cpus(dev):
    """Pin this rank to the CPU cores local to its GPU (NVML affinity) BEFORE the pinned host buffers are allocated,
    so that with several ranks on one box every rank's host copies stay on its GPU's NUMA node. Returns a short note
    for the JSON line; never fails the run."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(dev)
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = {64 * i + bit for i, w in enumerate(mask) for bit in range(64) if (w >> bit) & 1}
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return "gpu-local cpus not in this process's cpuset"
        os.sched_setaffinity(0, cpus)
        return "%d gpu-local cpus" % len(cpus)
    except Exception as e:  # NVML absent, CUDA_VISIBLE_DEVICES remapping, ...
        return "unbound (%s)" % type(e).__name__


def init_nccl(local_rank):
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local_rank)
    with stdout_to_stderr():
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        t = torch.zeros(1, device="cuda")
        dist.all_reduce(t)          # first collective: communicator set-up (and the banner) happen here
        torch.cuda.synchronize()
    return dist


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""

    def __init__(self, gpu):
        super().__init__(daemon=True)
        self.gpu, self.rows, self.stop_flag = gpu, [], False

    def run(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-i", str(self.gpu)],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            time.sleep(0.1)

    def keep_loaded(self, step, sync, iters):
        """For timed regions of a few ms (one nvidia-smi query takes longer): issue the SAME step `iters` more times,
        untimed, so that samples are taken under that load. `iters` must be the same on every rank (the long-horizon step
        exchanges data between ranks). Returns how the clocks were obtained (goes into `clocks`)."""
        if len(self.rows) >= 3:
            return "during the timed region"
        for _ in range(iters):
            step()
        sync()
        return (f"untimed continuation of the same step loop ({iters} steps) right after the timed region "
                "(the timed region is shorter than one nvidia-smi query)")

    def summary(self):
        self.stop_flag = True
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        sm = sorted(float(r[0]) for r in self.rows if r[0].replace(".", "").isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [nm for i, nm in enumerate(names) if any(r[2 + i].lower().startswith("active") for r in self.rows if len(r) >= 6)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": float(self.rows[0][1]), "reasons": reasons,
                "samples": len(self.rows)}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)), "measured"
    return {"hbm_gbs": 6650.0}, "fallback"


# ------------------------------------------------------------------------------------------
# CPU side: the reference's own ndlqr_Solve (oracle/_ref, unmodified sources) on the host cores
# ------------------------------------------------------------------------------------------
def cpu_reference_rate(n, m, N, knots, x0, budget_s=12.0, cores=None):
    """Solves/sec of the reference on this host: one problem per core, num_threads = 1 each
    (the strongest CPU configuration for a batch). Runs a bounded sample."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as O
    cores = cores or max(1, len(os.sched_getaffinity(0)))
    kind = "reference" if O.REF is not None else "port"
    nprob = knots.shape[0]
    done = [0] * cores
    t_end = [0.0]
    start_evt = threading.Event()

    def worker(tid):
        if kind == "reference":
            pp = [O.ref_problem_from_flat(n, m, N, knots[(tid + i * cores) % nprob], x0[(tid + i * cores) % nprob])
                  for i in range(2)]
            sp = O.REF.ndlqr_NewNdLqrSolver(n, m, N)
            O.REF.ndlqr_SetNumThreads(sp, 1)
        else:
            ps = O.PortSolver(n, m, N)
        start_evt.wait()
        i = 0
        while time.perf_counter() < t_end[0]:
            if kind == "reference":
                O.REF.ndlqr_ResetSolver(sp)
                O.REF.ndlqr_InitializeWithLQRProblem(pp[i % 2], sp)
                O.REF.ndlqr_Solve(sp)
            else:
                j = (tid + i * cores) % nprob
                ps.init(knots[j], x0[j])
                ps.solve(1)
            i += 1
            done[tid] = i

    th = [threading.Thread(target=worker, args=(t,)) for t in range(cores)]
    for t in th:
        t.start()
    time.sleep(0.5)  # setup
    t0 = time.perf_counter()
    t_end[0] = t0 + budget_s
    start_evt.set()
    for t in th:
        t.join()
    dt = time.perf_counter() - t0
    total = sum(done)
    return {"value": total / dt, "unit": "solves/s", "cores": cores, "kind": kind,
            "sample": f"{total} solves of the same n={n},m={m},N={N} problems in {dt:.1f}s, one problem per core, "
                      f"num_threads=1 per solve (ndlqr_Solve only + reset/init)"}


def cpu_riccati_baseline(n, m, N, knots, x0, budget_s=3.0):
    """The reference's serial Riccati solver (src/riccati_solve.c:7-24, ndlqr_SolveRiccati) on ONE host core: a timing
    baseline only -- its KKT residual is printed beside it because the recursion is unstable on this data."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import ctypes as C
    import oracle_lib as O
    zs = 2 * n + m
    if O.REF is not None:
        lib = O.REF
        lib.ndlqr_NewRiccatiSolver.restype = C.c_void_p
        lib.ndlqr_NewRiccatiSolver.argtypes = [C.c_void_p]
        lib.ndlqr_SolveRiccati.restype = C.c_int
        lib.ndlqr_SolveRiccati.argtypes = [C.c_void_p]
        lib.ndlqr_CopyRiccatiSolution.restype = C.c_int
        lib.ndlqr_CopyRiccatiSolution.argtypes = [C.c_void_p, C.c_void_p]
        lib.ndlqr_FreeRiccatiSolver.argtypes = [C.c_void_p]
        pp = O.ref_problem_from_flat(n, m, N, knots[0], x0[0])
        rs = lib.ndlqr_NewRiccatiSolver(C.cast(pp, C.c_void_p))
        lib.ndlqr_SolveRiccati(rs)
        cnt, t0 = 0, time.perf_counter()
        while time.perf_counter() - t0 < budget_s:
            lib.ndlqr_SolveRiccati(rs)
            cnt += 1
        dt = time.perf_counter() - t0
        z = np.zeros(N * zs)
        lib.ndlqr_CopyRiccatiSolution(rs, z.ctypes.data_as(C.c_void_p))
        lib.ndlqr_FreeRiccatiSolver(rs)
        kind = "reference"
    else:
        O.riccati(n, m, N, knots[0], x0[0])
        cnt, t0 = 0, time.perf_counter()
        while time.perf_counter() - t0 < budget_s:
            z = O.riccati(n, m, N, knots[0], x0[0])
            cnt += 1
        dt = time.perf_counter() - t0
        kind = "port"
    res = O.kkt_residual(n, m, N, knots[0], x0[0], z)
    return {"value": cnt / dt, "unit": "solves/s", "cores": 1, "kind": kind, "kkt_residual": res,
            "sample": f"{cnt} ndlqr_SolveRiccati calls on problem 0 in {dt:.1f}s, one core (timing baseline only)"}


def synth_only(n, m, N, seed, first, count):
    """Synthetic problems from rslqr_b200/libsynth.so (the generator alone, plain host C): the CPU reference arm uses
    this so that it never maps the product library (same generator, same bits as rslqr_b200.synth_problems)."""
    import ctypes as C
    lib = C.CDLL(os.path.join(ROOT, "rslqr_b200", "libsynth.so"))
    lib.rslqr_synth_fill.restype = C.c_int
    lib.rslqr_synth_fill.argtypes = [C.c_int, C.c_int, C.c_int, C.c_uint64, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    st = 2 * n + 2 * m + 1 + n * n + n * m + n
    knots, x0 = np.empty((count, N, st)), np.empty((count, n))
    if lib.rslqr_synth_fill(n, m, N, seed, first, count, knots.ctypes.data_as(C.c_void_p), x0.ctypes.data_as(C.c_void_p)) != 0:
        raise RuntimeError("rslqr_synth_fill failed")
    return knots, x0


def run_reference(args, rank, world):
    if rank != 0:
        return
    n, m, N = N_STATES, N_INPUTS, HORIZON
    cores = max(1, len(os.sched_getaffinity(0)))
    knots, x0 = synth_only(n, m, N, SEED, 0, min(args.batch, 4 * cores))
    per_step = max(2.0, min(20.0, 120.0 / max(1, args.steps + args.warmup)))
    for _ in range(args.warmup):
        cpu_reference_rate(n, m, N, knots, x0, budget_s=min(2.0, per_step), cores=cores)
    rates = [cpu_reference_rate(n, m, N, knots, x0, budget_s=per_step, cores=cores) for _ in range(args.steps)]
    val = float(np.mean([r["value"] for r in rates]))
    base = rates[-1]
    base["value"] = val
    try:
        base["riccati"] = cpu_riccati_baseline(n, m, N, knots[:1], x0[:1])
    except Exception as e:
        base["riccati"] = {"unavailable": repr(e)}
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": "solves/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * per_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": WORKLOAD, "batch_per_gpu": args.batch, "global_batch": args.batch * world, "seed": SEED,
                       "mode": "reference ndlqr_Solve on host cores", "parallelism": "one problem per host core, num_threads=1",
                       "note": "reference ndlqr_Solve (OpenMP/C, unmodified sources compiled by oracle/Makefile) on host cores; "
                               "bounded sample of the same generator and seed, extrapolated to the batch"},
            "native_libraries": ["oracle/_ref/libref_rslqr.so (or oracle/liboracle_port.so)", "rslqr_b200/libsynth.so (problem generator only)"],
            "cpu_baseline": base,
            "e2e": {"value": val, "unit": "solves/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------
# GPU side
# ------------------------------------------------------------------------------------------
def run_ours(args, rank, world, local_rank):
    import rslqr_b200 as R
    from rslqr_b200.lib import pinned_array
    n, m, N, B = args.n, args.m, HORIZON, args.batch
    dual = (n, m) != (12, 4) or bool(int(os.environ.get("RSLQR_DUAL", "0") or 0))
    hybrid = (n, m) == (12, 4) and not dual and not int(os.environ.get("RSLQR_NO_HYBRID", "0") or 0)
    dist = None
    cpu_binding = bind_to_gpu_cpus(local_rank) if world > 1 and not os.environ.get("RSLQR_BENCH_NO_BIND") else "not bound (one rank)"
    if world > 1:
        dist = init_nccl(local_rank)

    def barrier():
        if dist is not None:
            dist.barrier()

    def allmax(x):
        if dist is None:
            return x
        import torch
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    mode = {"auto": R.MODE_AUTO, "staged": R.MODE_STAGED, "fused": R.MODE_FUSED, "scan": R.MODE_SCAN}[args.mode]
    flags = R.KEEP_FACT | R.PROFILE if mode == R.MODE_STAGED else 0
    s = R.CudaSolver(n, m, N, batch=B, device=local_rank, flags=flags)
    if mode != R.MODE_AUTO:
        s.set_mode(mode)
    if s.mode == R.MODE_STAGED and not (flags & R.KEEP_FACT):
        s.close()
        flags = R.KEEP_FACT | R.PROFILE
        s = R.CudaSolver(n, m, N, batch=B, device=local_rank, flags=flags)
    staged = s.mode == R.MODE_STAGED
    st, zs = s.stride, s.zs
    # synthetic inputs, generated into pinned host memory (ids rank*B .. rank*B+B-1)
    h_knots = pinned_array((B, N, st))
    h_x0 = pinned_array((B, n))
    h_soln = pinned_array((B, N * zs))
    nthreads = min(16, max(1, len(os.sched_getaffinity(0))))
    chunks = np.array_split(np.arange(B), nthreads)

    def gen(ids):
        if len(ids):
            R.synth_problems(n, m, N, SEED, rank * B + int(ids[0]), len(ids), h_knots[ids[0]:ids[-1] + 1], h_x0[ids[0]:ids[-1] + 1])
    th = [threading.Thread(target=gen, args=(c,)) for c in chunks]
    [t.start() for t in th]
    [t.join() for t in th]
    s.upload(h_knots, h_x0)

    def step_resident():
        s.solve_async(True)    # (a staged solve clears its factor array itself, stream-ordered: no host sync in here)

    def step_e2e():
        # the user-facing call: host problems in, host solutions out (chunked, copies overlap the kernels)
        s.solve_host(h_knots, h_x0, h_soln, chunk=args.chunk)
        s.sync()

    for _ in range(args.warmup):
        step_resident()
        s.sync()
    launches_per_step = s.launches
    s.kernel_timing(enable=True, reset=True)
    sampler = ClockSampler(local_rank)
    sampler.start()
    barrier(); s.sync()
    s.mark(0)
    for _ in range(args.steps):
        step_resident()
    s.mark(1)
    ms_total = s.elapsed_ms(0, 1)
    s.sync(); barrier()
    ms_total = allmax(ms_total)
    k_ms, k_n = s.kernel_timing(enable=False, reset=True)
    prof = s.profile() if staged else None
    assert not s.flags().any(), "a Cholesky factorization failed"
    # end-to-end: host buffers in, host solutions out, copies inside the timed region
    for _ in range(min(2, args.warmup)):
        step_e2e()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_e2e()
    e2e_s = allmax(time.perf_counter() - t0)
    clocks = sampler.summary()
    barrier()

    ms_per_step = ms_total / args.steps
    value = world * B / (ms_per_step * 1e-3)
    pk, pk_kind = peaks()
    if staged:
        bytes_launch = schur_bytes_model(n, m, N, B)
        k_avg_ms = prof["schur"] / (s.K - 1) if prof and prof["schur"] > 0 else None
        ach = bytes_launch / (k_avg_ms * 1e-3) / 1e9 if k_avg_ms else None
        roof = {"kernel": "schur_kernel (staged, per level)", "bound": "hbm", "achieved": ach, "peak": pk["hbm_gbs"],
                "unit": "GB/s", "frac": ach / pk["hbm_gbs"] if ach else None, "traffic": None, "peak_kind": pk_kind,
                "algorithmic_bytes_per_launch": bytes_launch, "avg_launch_ms": k_avg_ms}
    elif s.mode == R.MODE_SCAN:
        # dominant kernel of the scan path: scan_leaf_kernel (levels 0 and 1 = 3/4 of all separators), one launch per
        # step, timed with CUDA events on the launching stream by the C-ABI (rslqr_cuda_kernel_timing)
        nsub = B * (N // 4)
        by_sub = scan_leaf_bytes_per_subtree(n, m, dual or hybrid)
        fl_sub = scan_leaf_flops_per_subtree(n, m)
        k_avg_ms = k_ms if k_n else ms_per_step
        ach = nsub * by_sub / (k_avg_ms * 1e-3) / 1e9
        fp64 = fp64_peak()
        roof = {"kernel": (f"dual_leaf_kernel<{n},{m}> (one warp per 4-knot subtree: leaf elements + levels 0..1 in dual form)" if dual else
                           f"scan_leaf_kernel<{n},{m}> (one warp per 4-knot subtree: leaf factors + levels 0..1 on boundary entries)"),
                "bound": "hbm", "achieved": ach, "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": ach / pk["hbm_gbs"],
                "traffic": SCAN_LEAF_TRAFFIC_BYTES_PER_PROBLEM * B if SCAN_LEAF_TRAFFIC_BYTES_PER_PROBLEM and (n, m, N) == (12, 4, 256) and not dual else None,
                "whole_step_hbm_frac": whole_step_bytes(n, m, N, dual or hybrid) * B / (ms_per_step * 1e-3) / 1e9 / pk["hbm_gbs"],
                "peak_kind": pk_kind, "algorithmic_bytes_per_launch": nsub * by_sub, "bytes_per_subtree": by_sub,
                "avg_launch_ms": k_avg_ms, "kernel_launches_timed": k_n, "kernel_share_of_step": k_avg_ms / ms_per_step,
                "binding_unit": "dependent chains of the 12 pivot steps at 12 warps/SM (LSU data pipe 74 % of peak, DRAM 22 % busy; "
                                "ncu --set full, profiles/r02_ncu_full.md)",
                "fp64_tflops_executed": nsub * fl_sub / (k_avg_ms * 1e-3) / 1e12, "fp64_frac_of_dmma_peak": nsub * fl_sub / (k_avg_ms * 1e-3) / 1e12 / fp64["tflops"],
                "flops_per_subtree": fl_sub,
                "reference_count_tflops_whole_step": flop_model(n, m, N) * B / (ms_per_step * 1e-3) / 1e12,
                "note": "achieved = algorithmic HBM bytes of the leaf kernel (4 flat knot blocks in, 1 boundary entry + 3 separator "
                        "records out per subtree, DESIGN.md section 5) / its CUDA-event time; its arithmetic intensity (3.7 flop/B) is "
                        "below the fp64 ridge (5.7 flop/B), so HBM is the bounding roofline; fp64_* report the executed FLOPs beside it"}
    else:
        # dominant kernel of the tiered path: tier0_kernel (leaves + the first log2(G) levels on chip), one launch
        # per step, timed with CUDA events on the launching stream by the C-ABI (rslqr_cuda_kernel_timing)
        G0 = int(os.environ.get("RSLQR_G0", "8"))
        G0 = 16 if G0 == 16 else 8
        fl_group = tier0_flops_per_group(n, m, G0)
        fl = fl_group * (N // G0) * B
        k_avg_ms = k_ms if k_n else ms_per_step
        ach = fl / (k_avg_ms * 1e-3) / 1e12
        fp64 = fp64_peak()
        roof = {"kernel": f"tier0_kernel<{n},{m},G={G0}> (leaves + levels 0..{G0.bit_length() - 2} on chip, DMMA Schur update)",
                "bound": "tensor", "achieved": ach, "peak": fp64["tflops"], "unit": "TFLOP/s", "frac": ach / fp64["tflops"],
                "traffic": TIER0_TRAFFIC_BYTES_PER_PROBLEM * B if (n, m, N, G0) == (12, 4, 256, 8) else None,
                "peak_kind": fp64["kind"],
                "flops_per_launch": fl, "flops_per_group": fl_group, "avg_launch_ms": k_avg_ms, "kernel_launches_timed": k_n,
                "kernel_share_of_step": k_avg_ms / ms_per_step,
                "reference_count_tflops_whole_step": flop_model(n, m, N) * B / (ms_per_step * 1e-3) / 1e12,
                "note": "achieved = FLOPs tier0 executes (DESIGN.md section 5) / its CUDA-event time; "
                        "reference_count = SURVEY 8d figure (32.2 MFLOP/solve as the reference executes it) over the whole step"}
    metric = METRIC if (n, m) == (12, 4) else f"LQR solves/sec (n={n},m={m},N={N} batched)"
    workload = WORKLOAD if (n, m) == (12, 4) else f"batched synthetic LQR n={n} m={m} N={N} (shape sweep, not a BASELINE config)"
    line = {"metric": metric, "value": value, "unit": "solves/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "value_kind": VALUE_KIND,
            "config": {"workload": workload, "batch_per_gpu": B, "global_batch": B * world, "seed": SEED,
                       "entry_form": "dual (Schur-complement elements)" if dual and s.mode == R.MODE_SCAN else "boundary rows",
                       "mode": {R.MODE_STAGED: "staged", R.MODE_FUSED: "fused (tiered)", R.MODE_SCAN: "scan"}[s.mode], "parallelism": f"batch-shard x{world}, no collective",
                       "l2": "inputs (%.2f GB/GPU) larger than L2; no explicit flush" % (B * N * st * 8 / 1e9),
                       "host_cpu_binding": cpu_binding},
            "clocks": clocks,
            "e2e": {"value": world * B * args.steps / e2e_s, "unit": "solves/s",
                    "h2d_bytes_per_step": int(h_knots.nbytes + h_x0.nbytes), "d2h_bytes_per_step": int(h_soln.nbytes)},
            "gpu_launches": int(launches_per_step * args.steps),
            "roofline": roof}
    if staged and prof:
        line["phase_ms"] = prof
    if rank == 0 and world == 1 and not args.no_cpu:
        line["cpu_baseline"] = cpu_reference_rate(n, m, N, h_knots[:64].copy(), h_x0[:64].copy(), budget_s=args.cpu_seconds)
        try:
            line["cpu_baseline"]["riccati"] = cpu_riccati_baseline(n, m, N, h_knots[:1].copy(), h_x0[:1].copy())
        except Exception as e:  # the baseline must never take the bench line down
            line["cpu_baseline"]["riccati"] = {"unavailable": repr(e)}
    if rank == 0 and world == 1 and (n, m) == (12, 4):
        try:
            line["riccati_gpu"] = gpu_riccati_comparison(R, s, B)
        except Exception as e:  # a comparison figure must never take the bench line down
            line["riccati_gpu"] = {"unavailable": repr(e)}
    s.close()
    if dist is not None and (n, m) == (12, 4) and not os.environ.get("RSLQR_BENCH_NO_LONG"):
        # SURVEY 8e row 2 on the same ranks (after every timed region of the batch workload)
        try:
            rec = long_horizon_record(R, dist, rank, world, local_rank)
            if rank == 0:
                line["long_horizon"] = rec
        except Exception as e:  # never take the headline line down; all ranks reach this point together or not at all
            if rank == 0:
                line["long_horizon"] = {"unavailable": repr(e)}
    if rank == 0:
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def kkt_residual_numpy(n, m, N, knots, x0, z):
    """2-norm of K z - b for the LQR KKT system, straight from the flat knot blocks; terminal-knot
    R, r, A, B, d are ignored like the reference does. A self-check of what bench.py reports, in numpy on the host."""
    Z = np.asarray(z).reshape(N, 2 * n + m)
    lam, x, u = Z[:, :n], Z[:, n:2 * n], Z[:, 2 * n:]
    oA = 2 * n + 2 * m + 1
    Q, Rr, q, r = knots[:, :n], knots[:, n:n + m], knots[:, n + m:2 * n + m], knots[:, 2 * n + m:2 * n + 2 * m]
    A = knots[:, oA:oA + n * n].reshape(N, n, n).transpose(0, 2, 1)            # column-major blocks
    B = knots[:, oA + n * n:oA + n * n + n * m].reshape(N, m, n).transpose(0, 2, 1)
    d = knots[:, oA + n * n + n * m:]
    res = [x[0] - x0]
    rx = Q * x - lam + q
    rx[:-1] += np.einsum("kij,ki->kj", A[:-1], lam[1:])
    res.append(rx)
    res.append(Rr[:-1] * u[:-1] + np.einsum("kij,ki->kj", B[:-1], lam[1:]) + r[:-1])
    res.append(np.einsum("kij,kj->ki", A[:-1], x[:-1]) + np.einsum("kij,kj->ki", B[:-1], u[:-1]) - x[1:] + d[:-1])
    return float(np.sqrt(sum(float(np.sum(np.square(t))) for t in res)))


def long_horizon_record(R, dist, rank, world, local_rank, steps=20, N=16384):
    """BASELINE configs[3] on the SAME ranks, after the timed batch region: one n=12, m=4, N=16384 problem split into
    `world` subtrees. Latency with CUDA events (max over ranks), the KKT residual of the gathered
    solution, whether it is bit-identical to the one-GPU solve of the same problem, and what the exchange moved."""
    import torch
    from rslqr_b200.dist import LongHorizon
    n, m = N_STATES, N_INPUTS
    knots, x0 = R.synth_problems(n, m, N, SEED, 0, 1)
    lh = LongHorizon(n, m, N, rank, world, local_rank, exchange="auto")
    lh.upload(knots[0], x0[0])
    for _ in range(5):
        lh.solve_async()
    lh.check()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(steps):
        lh.solve_async()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / steps], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    z = torch.from_numpy(lh.solution()).cuda()
    parts = [torch.empty_like(z) for _ in range(world)]
    dist.all_gather(parts, z)
    rec = None
    if rank == 0:
        zz = torch.cat(parts).cpu().numpy()
        one = R.CudaSolver(n, m, N, batch=1, device=local_rank, flags=0)
        one.upload(knots, x0)
        for _ in range(5):
            one.solve_async(True)
        one.sync()
        one.mark(6)
        for _ in range(steps):
            one.solve_async(True)
        one.mark(7)
        ms1 = one.elapsed_ms(6, 7) / steps
        z1 = one.solution()[0]
        launches1 = one.launches
        one.close()
        pay = lh.shard.payload_doubles * 8
        rec = {"workload": f"single long-horizon LQR n={n} m={m} N={N} (BASELINE configs[3])", "n_gpus": world,
               "latency_ms": ms, "latency_ms_one_gpu_same_box": ms1, "speedup_vs_one_gpu": ms1 / ms,
               "kkt_residual": kkt_residual_numpy(n, m, N, knots[0], x0[0], zz), "z_norm": float(np.linalg.norm(zz)),
               "bit_identical_to_1gpu": bool(np.array_equal(zz, z1)),
               "rel_diff_vs_1gpu": float(np.linalg.norm(zz - z1) / np.linalg.norm(z1)),
               "exchange": lh.exchange, "exchanges_per_solve": 1, "exchange_bytes_per_rank_per_solve": int(pay * (world - 1)),
               "kernel_launches_per_solve_per_rank": int(lh.shard.launches), "kernel_launches_per_solve_one_gpu": int(launches1),
               "timing": f"CUDA events on the launching stream around {steps} back-to-back solves, max over ranks"}
    lh.close()
    return rec


def gpu_riccati_comparison(R, s, B, reps=5):
    """the reference's own baseline algorithm (Riccati recursion, src/riccati_solve.c) on the same GPU and
    the same resident batch, one warp per problem -- NOT the path this repo accelerates; reported so that the
    nested-dissection number can be read against the best serial-in-time competitor. Outside every timed region above."""
    s.solve_async(True)
    s.sync()
    z_nd = s.solution(0, min(B, 64)).copy()
    for _ in range(2):
        s.riccati()
    s.mark(4)
    for _ in range(reps):
        R.LIB.rslqr_cuda_riccati_solve_async(s.h)
    s.mark(5)
    s.sync()
    ms = s.elapsed_ms(4, 5) / reps
    z_r = s.solution(0, min(B, 64))
    return {"value": B / (ms * 1e-3), "unit": "solves/s", "ms_per_step": ms, "kernels": "riccati_kernel + riccati_forward_kernel",
            "rel_diff_vs_nested_dissection": float(np.linalg.norm(z_r - z_nd) / np.linalg.norm(z_nd)),
            "note": "comparison only: Riccati recursion (serial in the horizon, parallel over the batch), same inputs resident in HBM"}


def run_long(args, rank, world, local_rank):
    """BASELINE configs[3]: ONE long-horizon problem (n=12, m=4, N=16384) split into `world` subtrees, one per GPU;
    the only exchange is one all-gather of the 7.2 KB per-shard payloads over NVLink (NCCL). Reports latency."""
    import torch
    import rslqr_b200 as R
    from rslqr_b200.dist import LongHorizon
    n, m, N = N_STATES, N_INPUTS, args.horizon
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        dist = init_nccl(local_rank)
    knots, x0 = R.synth_problems(n, m, N, SEED, 0, 1)
    lh = LongHorizon(n, m, N, rank, world, local_rank, exchange=args.exchange)
    lh.upload(knots[0], x0[0])
    for _ in range(max(args.warmup, 3)):
        lh.solve_async()
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(args.steps):
        lh.solve_async()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.steps
    if dist is not None:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    # end to end: upload of the shard from pinned host memory + solve + download of the shard's solution
    from rslqr_b200.lib import pinned_array
    Ns = N // world
    lo = rank * Ns
    h_kn = pinned_array((1, Ns, knots.shape[-1]))
    h_kn[0] = knots[0, lo:lo + Ns]
    h_head = pinned_array((1, n))
    h_head[0] = x0[0] if rank == 0 else knots[0, lo - 1, -n:]
    h_z = pinned_array((1, Ns * (2 * n + m)))
    for _ in range(2):
        lh.shard.upload(h_kn, h_head)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        lh.shard.upload(h_kn, h_head)
        lh.solve_async()
        torch.cuda.synchronize()
        z = lh.shard.solution(out=h_z)[0]
    e2e_ms = (time.perf_counter() - t0) / args.steps * 1e3
    if dist is not None:
        t = torch.tensor([e2e_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_ms = float(t.item())
    how = sampler.keep_loaded(lh.solve_async, torch.cuda.synchronize, int(min(20000, max(100, 600.0 / ms))))
    clocks = sampler.summary()
    clocks["sampled"] = how
    assert not lh.shard.flags().any() and np.isfinite(z).all()
    line = {"metric": f"single long-horizon LQR solve latency (n={n},m={m},N={N})", "value": ms, "unit": "ms", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms, "higher_is_better": False, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"single long-horizon LQR n={n} m={m} N={N} (BASELINE configs[3])", "seed": SEED,
                       "parallelism": f"{world} subtrees of {N // world} knots, one exchange of {lh.shard.payload_doubles * 8} B per rank",
                       "exchange": ("peer stores over NVLink into symmetric memory + arrival flags (own kernels)" if lh.exchange == "p2p"
                                    else "NCCL all-gather") if world > 1 else "none",
                       "l2": "working set (state 92 MB + inputs 31 MB over all ranks) vs 126 MB L2 per GPU; no explicit flush"},
            "clocks": clocks,
            "e2e": {"value": e2e_ms, "unit": "ms", "h2d_bytes_per_step": int(h_kn.nbytes + h_head.nbytes), "d2h_bytes_per_step": int(h_z.nbytes)},
            "gpu_launches": int(lh.shard.launches * args.steps)}
    if rank == 0 and args.cpu_long:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import oracle_lib as O
        cores = max(1, len(os.sched_getaffinity(0)))
        t0 = time.perf_counter()
        O.port_solve(n, m, N, knots[0], x0[0], threads=cores)
        line["cpu_baseline"] = {"value": (time.perf_counter() - t0) * 1e3, "unit": "ms", "cores": cores, "kind": "port",
                                "sample": "one solve of the same problem, oracle C port with OpenMP over all host cores (includes init)"}
    if rank == 0:
        print(json.dumps(line), flush=True)
    lh.close()
    if dist is not None:
        dist.destroy_process_group()


def big_flops_per_solve(n, m, N):
    """FLOPs the large-block dual path (kernels_big.cuh) executes per solve: per merge the LDL^T (n^3/6 FMA), the forward
    substitution with 2n+1 right-hand sides ((2n+1) n^2 / 2 FMA) and the Schur update (Sab': n^3, sa', sb': 2 n^2,
    Saa' and Sbb': the lower block triangle of 8 x 8 tiles only, n^3 (t+1)/(2t) each with t = n/8), per knot the leaf
    block Sbb (n^2 (n+m) (t+1)/(2t) FMA)."""
    t = n // 8
    tri = (t + 1) / (2.0 * t)
    merge = n ** 3 / 6 + (2 * n + 1) * n * n / 2 + n ** 3 + 2 * n * n + 2 * tri * n ** 3
    return 2.0 * ((N - 1) * merge + N * n * n * (n + m) * tri)


def run_large(args, rank, world, local_rank):
    """BASELINE configs[4]: larger-state LQR n=64 m=16 N=1024, batch 512 per run (split over ranks).
    Default: the dual-form path for large blocks (kernels_big.cuh: elements in HBM, a CTA per merge; 0.22 GB per problem).
    --mode staged: the per-level staged path with the dense factor array (755 MB per problem, chunks of 16)."""
    import rslqr_b200 as R
    n, m, N = 64, 16, 1024
    total = args.batch if args.batch != 4096 else 512
    staged = args.mode == "staged"
    share = total // world
    chunk = min(args.chunk or (16 if staged else 128), share)
    dist = init_nccl(local_rank) if world > 1 else None
    s = R.CudaSolver(n, m, N, batch=chunk, device=local_rank, flags=R.KEEP_FACT if staged else 0)
    if staged:
        s.set_mode(R.MODE_STAGED)
    kn, x0 = R.synth_problems(n, m, N, SEED, rank * share, chunk)
    s.upload(kn, x0)

    def one_chunk():
        s.solve_async(True)
    for _ in range(max(1, args.warmup // 3)):
        one_chunk()
    s.sync()
    nchunks = share // chunk
    steps = max(1, min(args.steps, 2))
    sampler = ClockSampler(local_rank)
    sampler.start()
    if dist is not None:
        dist.barrier()
    s.mark(0)
    for _ in range(steps * nchunks):   # every chunk re-solves resident problems of the same shape (inputs 44 MB each)
        one_chunk()
    s.mark(1)
    ms = s.elapsed_ms(0, 1) / steps
    s.sync()
    if dist is not None:
        import torch
        tt = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms = float(tt.item())
    assert not s.flags().any()
    launches = s.launches
    value = total / (ms * 1e-3)
    fp64 = fp64_peak()
    line = {"metric": "LQR solves/sec (n=64,m=16,N=1024 batched)", "value": value, "unit": "solves/s", "n_gpus": world,
            "steps": steps, "warmup": max(1, args.warmup // 3), "ms_per_step": ms, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "larger-state LQR n=64 m=16 N=1024 batch %d (BASELINE configs[4])" % total, "chunk": chunk,
                       "mode": "staged (DMMA products / Schur, dense factor array)" if staged else "scan, dual form for large blocks (kernels_big.cuh)",
                       "l2": "%s per problem, larger than L2" % ("factor array 755 MB" if staged else "elements + records 220 MB")},
            "clocks": sampler.summary(), "gpu_launches": int(launches * nchunks * steps)}
    ref_tf = flop_model(n, m, N) * value / 1e12
    if staged:
        line["roofline"] = {"kernel": "whole step (schur_mma_kernel dominates)", "bound": "tensor", "achieved": ref_tf, "peak": fp64["tflops"],
                            "unit": "TFLOP/s", "frac": ref_tf / fp64["tflops"], "traffic": None, "peak_kind": fp64["kind"],
                            "note": "structurally non-zero FLOPs of the reference algorithm (SURVEY 8d: 23.6 GFLOP per solve) / step time"}
        # factor once, solve many: the chunk just solved keeps its factorization; time the rhs phase alone
        reps = 5
        R.LIB.rslqr_cuda_solve_rhs_async(s.h, 1)
        s.sync()
        s.mark(2)
        for _ in range(reps):
            R.LIB.rslqr_cuda_solve_rhs_async(s.h, 1)
        s.mark(3)
        s.sync()
        ms_rhs = s.elapsed_ms(2, 3) / reps
        line["solve_many"] = {"value": chunk / (ms_rhs * 1e-3) * world, "unit": "rhs solves/s", "ms_per_chunk": ms_rhs, "chunk": chunk,
                              "full_solve_ms_per_chunk": ms / nchunks,
                              "note": "rhs phase only (pack + leaf scaling + K rhs levels) on the stored staged factorization; per rank x world"}
    else:
        tf = big_flops_per_solve(n, m, N) * value / 1e12
        line["roofline"] = {"kernel": "whole step (big_leaf_kernel, then big_factor_reg_kernel + big_solve_kernel + big_schur_kernel per level; profiles/r02_ncu_big.md)", "bound": "tensor", "achieved": tf,
                            "peak": fp64["tflops"], "unit": "TFLOP/s", "frac": tf / fp64["tflops"], "traffic": None, "peak_kind": fp64["kind"],
                            "flops_per_solve_executed": big_flops_per_solve(n, m, N),
                            "reference_count_tflops": ref_tf, "reference_count_frac": ref_tf / fp64["tflops"],
                            "note": "achieved = FLOPs this path executes (2.2 GFLOP per solve: interior knots are not updated above the "
                                    "leaves) / step time; reference_count = the reference's structurally non-zero 23.6 GFLOP per solve "
                                    "over the same time (can exceed the peak: the work is not done)"}
    if rank == 0:
        print(json.dumps(line), flush=True)
    s.close()
    if dist is not None:
        dist.destroy_process_group()


def run_single(args, rank, world, local_rank):
    """BASELINE configs[1]: the shipped lqr_prob_256.json (n=6, m=3, N=256), ONE solve through the reference-named C host
    API (ndlqr_InitializeWithLQRProblem + ndlqr_Solve + solution copy), beside the reference on the host cores."""
    if rank != 0:
        return
    import ctypes as C
    import rslqr_b200 as R
    path = os.path.join(ROOT, "tests", "golden", "lqr_prob_256.json")
    LIB = R.LIB
    pp = LIB.ndlqr_ReadLQRProblemJSONFile(path.encode())
    n, m, N = 6, 3, 256
    sv = R.NdLqrSolver(n, m, N)
    z = np.zeros(N * (2 * n + m))
    def once():
        sv.reset(); sv.initialize(pp); sv.solve()
        LIB.ndlqr_CopySolution(sv.p, z.ctypes.data_as(C.POINTER(C.c_double)))
    for _ in range(max(3, args.warmup)):
        once()
    sampler = ClockSampler(0)
    sampler.start()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        once()
    e2e_ms = (time.perf_counter() - t0) / args.steps * 1e3
    dev_ms = float(sv.c.solve_time_ms)
    launches_default = int(LIB.rslqr_cuda_kernel_launches(sv.c.device))
    # solver->mirror_fact = 1 (or RSLQR_MIRROR_FACT=1): every solve runs the reference's phases one launch each and
    # copies the dense `fact` array (1.5 MB here) back into solver->fact, which is what the reference leaves there
    sv.c.mirror_fact = 1
    for _ in range(5):
        once()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        once()
    mirror_e2e_ms = (time.perf_counter() - t0) / args.steps * 1e3
    mirror_ms = float(sv.c.solve_time_ms)
    launches_mirror = int(LIB.rslqr_cuda_kernel_launches(sv.c.device))
    sv.c.mirror_fact = 0
    how = sampler.keep_loaded(once, lambda: None, int(min(5000, max(100, 600.0 / max(e2e_ms, 1e-3)))))
    clocks = sampler.summary()
    clocks["sampled"] = how
    want = np.array(json.load(open(path))["soln"]).reshape(-1)
    err = float(np.linalg.norm(z[:want.size] - want))
    line = {"metric": "single-solve latency (lqr_prob_256.json n=6,m=3,N=256)", "value": dev_ms, "unit": "ms", "n_gpus": 1,
            "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": dev_ms, "higher_is_better": False, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "shipped fixture",
            "config": {"workload": "lqr_prob_256.json single solve via ndlqr_Solve (BASELINE configs[1])", "abs_err_vs_json_soln": err,
                       "ndlqr_Solve_ms_with_fact_mirror (solver->mirror_fact = 1)": mirror_ms,
                       "reset+init+solve+copy_ms_with_fact_mirror": mirror_e2e_ms,
                       "note": "value = ndlqr_Solve as a consumer calls it (solve_time_ms: rhs upload + on-chip kernels + solution "
                               "download); solver->fact is refreshed on request (ndlqr_SyncFactMirror) or by every solve with mirror_fact = 1"},
            "clocks": clocks,
            "e2e": {"value": e2e_ms, "unit": "ms", "h2d_bytes_per_step": int(N * R.knot_stride(n, m) * 8), "d2h_bytes_per_step": int(z.nbytes)},
            "gpu_launches": launches_default * args.steps, "gpu_launches_per_solve": launches_default,
            "gpu_launches_per_solve_with_fact_mirror": launches_mirror}
    if not args.no_cpu:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import oracle_lib as O
        from rslqr_b200.lib import problem_to_flat
        kn, x0 = problem_to_flat(pp, n, m)
        cores = max(1, len(os.sched_getaffinity(0)))
        base = {"unit": "ms", "kind": "reference" if O.REF is not None else "port", "cores": cores}
        if O.REF is not None:
            for th in sorted({1, max(1, cores // 2), cores}):
                best = min(O.ref_solve(n, m, N, kn, x0, threads=th, want_fact=False)[2] for _ in range(6))
                base["ndlqr_ms_threads_%d" % th] = best
            base["value"] = min(v for k, v in base.items() if k.startswith("ndlqr_ms"))
        r = cpu_riccati_baseline(n, m, N, kn[None], x0[None], budget_s=1.0)
        base["riccati_ms"] = 1e3 / r["value"]; base["riccati_kkt_residual"] = r["kkt_residual"]
        base["sample"] = "best of 6 ndlqr_Solve calls per thread count (solve_time_ms), Riccati averaged over 1 s"
        line["cpu_baseline"] = base
    print(json.dumps(line), flush=True)
    sv.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=4096)
    ap.add_argument("--n", type=int, default=N_STATES, help="states (batch workload; default = the BASELINE shape)")
    ap.add_argument("--m", type=int, default=N_INPUTS, help="inputs (batch workload)")
    ap.add_argument("--mode", default="auto", choices=["auto", "staged", "fused", "scan"])
    ap.add_argument("--workload", default="batch", choices=["batch", "long", "large", "single"],
                    help="batch = BASELINE configs[2] (default, the headline metric); long = configs[3] single long horizon")
    ap.add_argument("--horizon", type=int, default=16384, help="horizon of the long workload")
    ap.add_argument("--exchange", default="auto", choices=["auto", "p2p", "nccl"],
                    help="long workload: transport of the shard payloads (auto = p2p when symmetric memory works)")
    ap.add_argument("--cpu-long", action="store_true", help="long workload: also time one CPU solve (takes ~10-60 s)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--chunk", type=int, default=0, help="problems per pipeline chunk of the e2e call (0 = automatic)")
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
    elif args.workload == "long":
        run_long(args, rank, world, local_rank)
    elif args.workload == "large":
        run_large(args, rank, world, local_rank)
    elif args.workload == "single":
        run_single(args, rank, world, local_rank)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
