#!/usr/bin/env python
"""Benchmark of the PPLS_simult EM loop: FP64 EM iterations/s and fraction of the HBM roofline.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c3|c2|...]

A "step" is one EM iteration = one pass of the streaming kernel over the resident data plus the tiny
update.  Headline workload = BASELINE.json's headline shape (very tall N=1e7, p=1000, q=200, r=3; 96 GB,
resident in one B200's HBM).  With N GPUs the SAME total shape is row-sharded (strong scaling), one process per
GPU (torchrun); per iteration the ranks exchange one (p+q)r+3r^2-vector -- by one-shot peer stores over NVLink
summed in rank order inside the update (default) or by NCCL (PPLS_XCHG=0).  The other named shapes of
BASELINE.json (C2 tall, C4 wide column-sharded, C5 batched multi-start) ride in the same JSON line under
"other_workloads", and at N > 1 an untimed "parity" block checks the sharded fit against the CPU oracle.
Prints ONE JSON line on rank 0 and leaves through the interpreter's normal exit.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

# stdout is for the ONE JSON line.  Native libraries print there too (NCCL's "NCCL version ..." line under NCCL_DEBUG=VERSION,
# which ignores NCCL_DEBUG_FILE), so the real stdout is kept aside for emit() and file descriptor 1 is pointed at stderr.
sys.stdout.flush()
_JSON_FD = os.dup(1)
os.dup2(2, 1)


def emit(obj):
    os.write(_JSON_FD, (json.dumps(obj) + "\n").encode())


import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (N, p, q, r)   -- BASELINE.json configs
    "c1": (500, 20, 10, 2),
    "c2": (1_000_000, 500, 100, 3),
    "c4": (2000, 50000, 1000, 5),
    "c3": (10_000_000, 1000, 200, 3),
    "c5": (100_000, 200, 50, 2),          # x 64 starts (batched multi-start)
}
C5_STARTS = 64
METRIC = "FP64 EM iterations/sec at N x p x q x r; achieved HBM GB/s vs peak"
UNIT = "EM iterations/s"
SEED = 20261106


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, burst copy)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def workload_label(name, N, p, q, r):
    return f"{name}: N={N} p={p} q={q} r={r}" + (f" x {C5_STARTS} starts" if name == "c5" else "")


class ClockSampler:
    """Samples SM clocks and throttle reasons during the timed region (pynvml, 20 ms)."""

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz, self._stop = [], set(), None, threading.Event()
        self.index = index
        self._t = None
        self.active = False
        self.nv = None

    def prepare(self):
        """NVML set-up and the sampling thread, BEFORE the barrier that opens the timed region (NVML
        initialisation takes milliseconds and would skew the ranks); samples are kept once `start()` ran."""
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None
            return
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()

    def start(self):
        self.active = True

    def _run(self):
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
        }
        while not self._stop.is_set():
            if not self.active:
                self._stop.wait(0.002)
                continue
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.02)

    def stop(self):
        self._stop.set()
        if self._t:
            self._t.join(timeout=2)
        if self.nv is not None:
            try:
                self.nv.nvmlShutdown()
            except Exception:
                pass
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


def host_budget_bytes():
    """Host memory this process may still take: MemAvailable, capped by the cgroup limit when there is one."""
    avail = 0
    try:
        for line in open("/proc/meminfo"):
            if line.startswith("MemAvailable:"):
                avail = int(line.split()[1]) * 1024
                break
    except Exception:
        return 0
    for lim, cur in (("/sys/fs/cgroup/memory.max", "/sys/fs/cgroup/memory.current"),
                     ("/sys/fs/cgroup/memory/memory.limit_in_bytes", "/sys/fs/cgroup/memory/memory.usage_in_bytes")):
        try:
            v = open(lim).read().strip()
            if v != "max":
                avail = min(avail, int(v) - int(open(cur).read().strip()))
            break
        except Exception:
            continue
    return max(avail, 0)


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def set_cpu_threads(oracle_lib=None):
    """Give the CPU legs every host core and report what they really use: torchrun exports OMP_NUM_THREADS=1 to
    its workers, and numpy's BLAS reads that at import.  Returns (threads numpy's BLAS uses, OpenMP threads of the
    fused C restatement)."""
    n = host_cores()
    blas = 1
    try:
        from threadpoolctl import threadpool_info, threadpool_limits
        threadpool_limits(limits=n)
        pools = [p for p in threadpool_info() if p.get("user_api") == "blas"] or threadpool_info()
        blas = max([int(p.get("num_threads", 1)) for p in pools] or [1])
    except Exception:
        pass
    omp = 1
    if oracle_lib is not None:
        oracle_lib.ppls_oracle_set_threads(n)
        omp = int(oracle_lib.ppls_oracle_num_threads())
    return blas, omp


class PinnedPair:
    """Column-major host X (n x p) and Y (n x q), page-locked with cudaHostRegister (torch's pinned allocator
    rounds sizes up to a power of two, too much at 96 GB)."""

    def __init__(self, n, p, q):
        import torch
        self.rt = torch.cuda.cudart()
        self.X = np.empty((n, p), order="F")
        self.Y = np.empty((n, q), order="F")
        self.reg = []
        for a in (self.X, self.Y):
            rc = self.rt.cudaHostRegister(a.ctypes.data, a.nbytes, 0)
            if int(rc) != 0:
                self.close()
                raise RuntimeError(f"cudaHostRegister of {a.nbytes} bytes failed: {rc}")
            self.reg.append(a.ctypes.data)

    def close(self):
        for ptr in self.reg:
            self.rt.cudaHostUnregister(ptr)
        self.reg = []
        self.X = self.Y = None


def shard_rows(N, rank, world):
    lo, hi = N * rank // world, N * (rank + 1) // world
    return lo, hi - lo


class Env:
    """Process-level state of one rank: torch.distributed for the plumbing, the library context for the work."""

    def __init__(self, args):
        import torch
        import torch.distributed as dist
        from ppls_b200 import Context
        self.torch, self.dist = torch, dist
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if self.world != args.gpus and self.world > 1:
            raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={self.world}")
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        if self.world > 1:
            dist.init_process_group("nccl", device_id=self.dev)
            idt = torch.zeros(128, dtype=torch.uint8, device=self.dev)
            if self.rank == 0:
                idt.copy_(torch.frombuffer(bytearray(Context.nccl_unique_id()), dtype=torch.uint8))
            dist.broadcast(idt, 0)
            self.ctx = Context(self.local, self.rank, self.world, bytes(idt.cpu().numpy().tobytes()))
            # host-side barrier (gloo): a rank that waits here runs NO kernel on its GPU (an NCCL barrier would spin one)
            self.host_group = dist.new_group(backend="gloo")
        else:
            self.ctx = Context(self.local)

    def barrier(self):
        self.ctx.sync()
        self.torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()
            self.torch.cuda.synchronize()

    def maxf(self, *vals):
        if self.world == 1:
            return [float(v) for v in vals]
        t = self.torch.tensor(list(vals), dtype=self.torch.float64, device=self.dev)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return [float(x) for x in t.cpu()]

    def close(self):
        """Orderly teardown on EVERY rank (communicators are collective objects), then a normal return: the
        interpreter's own exit runs, so the driver's exit hooks see which native libraries were loaded."""
        self.barrier()
        self.ctx.close()
        if self.world > 1:
            self.dist.barrier()
            self.dist.destroy_process_group()


def parity_block(env):
    """Untimed, N > 1 only: the same ranks fit a 20001 x 301 x 77, r = 3, 12-step problem (what tests/mgpu_rowshard.py
    does; the driver's 1-GPU test run skips that test) and rank 0 compares it with the CPU oracle -- the checker,
    not the thing measured."""
    from ppls_b200 import Data, Fit, synth
    torch, dist = env.torch, env.dist
    N, p, q, r, steps = 20001, 301, 77, 3, 12
    truth = synth.truth_params(p, q, r, 0.5, 0.5, 0.3)
    st = synth.start_params(truth, 3)
    lo, n_local = shard_rows(N, env.rank, env.world)
    d = Data.synthetic(env.ctx, n_local, p, q, truth, seed=4242, row0=lo)
    f = Fit(d, st, r, steps, -np.inf, "SVD")
    f.run()
    res = f.result(want_mu=False)
    f.close(); d.close()
    t = torch.from_numpy(np.concatenate([res["W"].ravel(), res["C"].ravel(), res["loglik"]])).to(env.dev)
    ts = [torch.empty_like(t) for _ in range(env.world)]
    dist.all_gather(ts, t)
    same = all(torch.equal(ts[0], x) for x in ts)
    out = None
    if env.rank == 0:
        from oracle import cbind, ppls_fused as F
        Lo = cbind.load()
        X, Y = cbind.gen_data(Lo, 4242, 0, N, truth)
        o = F.fused_loop(Lo, X, Y, st["W"], st["C"], st["B"], st["sigE"], st["sigF"], st["sigH"], st["sigT"],
                         EMsteps=steps, atol=-np.inf, want_mu=False)
        rel = lambda a, b: float(np.max(np.abs(np.asarray(a) - np.asarray(b))) / np.max(np.abs(np.asarray(b))))   # noqa: E731
        out = {"shape": f"N={N} p={p} q={q} r={r}, {steps} EM steps, row-sharded over {env.world} ranks",
               "loglik_rel": rel(res["loglik"], o["loglik"]), "W_rel": rel(res["W"], o["W"]), "C_rel": rel(res["C"], o["C"]),
               "B_rel": rel(res["B"], o["B"]), "bit_identical_ranks": bool(same),
               "pass": bool(same and rel(res["loglik"], o["loglik"]) < 1e-10 and rel(res["W"], o["W"]) < 1e-8),
               "checker": "oracle/ppls_fused.py (fused CPU restatement, rank 0)"}
    env.barrier()
    return out


def bench_stream(env, name, K, W, args, want_e2e):
    """C1-C4: K timed EM iterations on the resident (sharded) data; optionally the end-to-end call with host buffers."""
    from ppls_b200 import Data, Fit, synth
    torch, world, rank = env.torch, env.world, env.rank
    ctx, dev = env.ctx, env.dev
    N, p, q, r = WORKLOADS[name.split("_")[0]]          # "c4_rows": C4's shape, row-sharded (bench-only comparison)
    if args.rows and name == args.workload:
        N = args.rows
    free_b, _ = torch.cuda.mem_get_info(dev)
    row0, n_local = shard_rows(N, rank, world)
    # c4 (wide omics) is COLUMN-sharded over the ranks; every other workload is row-sharded
    colmode = name == "c4" and world > 1
    need = 8 * (N * (p // world + q // world + 8) if colmode else n_local * (p + q + 4)) + (2 << 30)
    if need > free_b:
        raise SystemExit(f"workload {name} needs {need/2**30:.1f} GiB per GPU, {free_b/2**30:.1f} free")
    truth = synth.truth_params(p, q, r)
    start = synth.start_params(truth, seed=1)
    t0 = time.perf_counter()
    if colmode:
        cx, cy = slice(p * rank // world, p * (rank + 1) // world), slice(q * rank // world, q * (rank + 1) // world)
        n_local = N
        data = Data.synthetic_cols(ctx, N, truth, SEED, cx, cy)
        start_l = dict(start, W=start["W"][cx], C=start["C"][cy])
    else:
        data = Data.synthetic(ctx, n_local, p, q, truth, seed=SEED, row0=row0)
        start_l = start
    gen_s = time.perf_counter() - t0

    fit = Fit(data, start_l, r, EMsteps=W + K + 8, atol=-float("inf"), type="SVD", timing=True, check_every=W + K + 8)
    if W:
        fit.iterate(W)
    tm0 = fit.timing()
    sampler = ClockSampler(env.local)
    sampler.prepare()
    env.barrier()
    sampler.start()
    h0 = time.perf_counter()
    done, fin = fit.iterate(K)
    env.barrier()
    wall_ms = (time.perf_counter() - h0) * 1e3
    clocks = sampler.stop()
    tm1 = fit.timing()
    assert done == W + K and not fin, (done, fin)
    loop_ms = tm1["loop_ms"] - tm0["loop_ms"]
    k1_ms = (tm1["stream_ms"] - tm0["stream_ms"]) / max(tm1["stream_launches"] - tm0["stream_launches"], 1)
    tail_ms = (tm1["tail_ms"] - tm0["tail_ms"]) / K
    scalar_ms = (tm1.get("scalar_ms", 0.0) - tm0.get("scalar_ms", 0.0)) / K
    launches = tm1["kernel_launches"] - tm0["kernel_launches"]
    per_rank = None
    if world > 1:
        t = torch.tensor([loop_ms, k1_ms, wall_ms, tail_ms], dtype=torch.float64, device=dev)
        ts = [torch.empty_like(t) for _ in range(world)]
        env.dist.all_gather(ts, t)
        per_rank = [[round(float(x), 4) for x in (u[0] / K, u[1], u[3])] for u in (v.cpu() for v in ts)]
        loop_ms, k1_ms, wall_ms = env.maxf(loop_ms, k1_ms, wall_ms)
    fit.close()
    read_probe = None
    if not colmode and data.n * (data.p + data.q) * 8 >= (1 << 30):
        read_probe = ctx.probe("hbm_read", data)            # read-only ceiling of THIS GPU over the same resident bytes

    # ---- e2e: the public call with HOST buffers (upload + K steps + download) ----
    # Full shard when the host can hold it pinned (the device copy is freed first: two copies of C3 exceed
    # the HBM), else a row sample scaled to the shard; `e2e.sample` says which.
    e2e = None
    if want_e2e:
        shard_bytes = 8 * n_local * (data.p + data.q)
        full = (not colmode and args.e2e_rows == 0 and shard_bytes > 8 * 1_000_000 * (data.p + data.q)
                and shard_bytes * world <= 0.55 * host_budget_bytes())
        if world > 1:                                    # every rank takes the same branch
            t = torch.tensor([1.0 if full else 0.0], dtype=torch.float64, device=dev)
            env.dist.all_reduce(t, op=env.dist.ReduceOp.MIN)
            full = bool(t.cpu()[0] > 0.5)
        ns = n_local if full else min(n_local, args.e2e_rows or 1_000_000)
        pinned = None
        if full:
            try:
                pinned = PinnedPair(ns, data.p, data.q)
            except Exception as ex:                      # cannot pin that much here: fall back to the sample
                print(f"[bench] full-size pinned buffers unavailable ({ex}); using a row sample", file=sys.stderr)
                full, ns = False, min(n_local, 1_000_000)
        if pinned is not None:
            Xn, Yn = pinned.X, pinned.Y
            data.export(0, ns, out=(Xn, Yn))             # product API; fills the host-side inputs
            data.close()                                 # the e2e call allocates its own device copy
        else:
            Xp = torch.empty((data.p, ns), dtype=torch.float64, pin_memory=True)
            Yp = torch.empty((data.q, ns), dtype=torch.float64, pin_memory=True)
            Xn, Yn = Xp.numpy().T, Yp.numpy().T           # column-major views (R layout), pinned
            data.export(0, ns, out=(Xn, Yn))
        env.barrier()
        e0 = time.perf_counter()
        d2 = Data.from_host_cols(ctx, Xn, Yn) if colmode else Data.from_host(ctx, Xn, Yn)
        up_s = time.perf_counter() - e0                  # from_host returns after the upload has completed
        f2 = Fit(d2, start_l, r, EMsteps=K, atol=-float("inf"), type="SVD")
        f2.run()
        res = f2.result(want_mu=False)
        ctx.sync()
        e_s = time.perf_counter() - e0
        f2.close(); d2.close()
        (e_s,) = env.maxf(e_s)
        scale = n_local / ns
        what = (f"the full shard, {ns} rows/GPU" if ns == n_local else f"{ns} rows/GPU")
        e2e = {"value": K / (e_s * scale), "unit": UNIT, "h2d_bytes_per_step": int(8 * ns * (data.p + data.q) * scale / K),
               "d2h_bytes_per_step": int(8 * ((p + q) * r + 2 * r + 3 + K) / K),
               "sample": f"one PPLS_simult-style call (upload {what} from pinned host memory, {K} EM steps, "
                         f"download estimates+loglik)" + ("" if ns == n_local else f", time scaled x{scale:.1f} to the full shard"),
               "full_size": ns == n_local,
               "seconds_sample": e_s, "upload_seconds_sample": up_s,
               "upload_gbs": 8 * ns * (data.p + data.q) / up_s / 1e9, "loglik_last": float(res["loglik"][-1])}
        if pinned is not None:
            pinned.close()
        else:
            del Xp, Yp
    data.close()

    peak, peak_src = measured_peak()
    bytes_pass = int(tm1["bytes_per_pass"])              # algorithmic bytes per EM iteration on this rank:
    #   8*n*(p+q) for the fused kernel, twice that for the two-pass wide-row path
    wide = bytes_pass > 8 * n_local * (data.p + data.q)
    achieved = bytes_pass / (k1_ms * 1e-3) / 1e9
    traffic = None                                       # dram bytes per launch, from the committed ncu capture
    tpath = os.path.join(ROOT, "profiles", "k1_traffic.json")
    if os.path.exists(tpath) and not wide:
        try:
            tj = json.load(open(tpath))
            if tj.get("workload") == name:
                traffic = tj["dram_bytes_per_row"] * n_local
        except Exception:
            pass
    return {
        "value": K / (loop_ms * 1e-3), "unit": UNIT, "ms_per_step": loop_ms / K, "scaling": "strong",
        "config": {"workload": workload_label(name, N, p, q, r),
                   "sharding": (f"column-sharded over {world} ranks" if colmode else f"row-sharded {world} x {n_local}"),
                   "l2": "inputs larger than L2 (no flush needed)" if bytes_pass > (256 << 20) else "inputs fit L2: L2-resident between iterations",
                   "generate_s": round(gen_s, 2), "wall_ms_per_step": wall_ms / K,
                   "exchange": ("none (1 GPU)" if world == 1 else "one-shot peer stores + flags, summed in rank order inside the update kernel"
                                if tm1.get("peer_exchange") else "NCCL allreduce"),
                   "tail_ms_per_step": tail_ms, "tail_scalar_part_ms_per_step": scalar_ms,
                   "per_rank_ms_step_k1_tail": per_rank},
        "clocks": clocks, "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic, "peak_source": peak_src,
                     "kernel": ("wide_kernel x2 (project + accumulate passes; algorithmic bytes count both)" if wide else
                                "wide3_kernel (single pass from HBM, second read from L2; algorithmic bytes count one)"
                                if name.startswith("c4") else "em_stream_kernel"),
                     "kernel_ms": k1_ms, "algorithmic_bytes_per_launch": bytes_pass,
                     "frac_of_nominal_8000": achieved / 8000.0,
                     "read_only_probe_gbs": read_probe},
        "e2e": e2e,
    }


def bench_c5(env, K, W, args):
    """BASELINE config 5: 64 start states of N=1e5, p=200, q=50, r=2 over one resident dataset; value = batch EM
    iterations/s (one iteration = all 64 starts advance one step), time = max over ranks.  At N > 1 the ROWS are split
    (every rank runs all 64 starts on its rows; one allreduce of the starts' moment vectors per iteration): splitting the
    STARTS instead (SURVEY.md 8e's other option, no collective) does not scale with this kernel -- its GEMM tile is 128
    components wide, and 32 starts x 2 components fill half of it (measured on 2 GPUs: 1677 vs 1729 batch-it/s on one)."""
    from ppls_b200 import Data, PPLS_simult_batched, synth
    world, rank = env.world, env.rank
    ctx = env.ctx
    N, p, q, r = WORKLOADS["c5"]
    if args.rows and args.workload == "c5":
        N = args.rows
    truth = synth.truth_params(p, q, r)
    row0, n_local = shard_rows(N, rank, world)
    data = Data.synthetic(ctx, n_local, p, q, truth, seed=SEED + 4, row0=row0)
    lo, hi = 0, C5_STARTS
    inits = [synth.start_params(truth, seed=1000 + s) for s in range(lo, hi)]
    W = max(W, 1)
    PPLS_simult_batched(None, None, r, inits, EMsteps=W, atol=-float("inf"), data=data)
    sampler = ClockSampler(env.local)
    sampler.prepare()
    ctx.sync()
    env.barrier()
    sampler.start()
    h0 = time.perf_counter()
    res, tm = PPLS_simult_batched(None, None, r, inits, EMsteps=K, atol=-float("inf"), data=data, timing=True)
    wall_ms = (time.perf_counter() - h0) * 1e3
    clocks = sampler.stop()
    (loop_ms,) = env.maxf(tm["loop_ms"])
    fp64_peak = ctx.probe("fp64_dmma")
    flops = C5_STARTS * 4.0 * N * (p + q) * r * tm["passes"]
    best = float(max(x["loglik"][-1] for x in res))
    data.close()
    return {"value": K / (loop_ms * 1e-3), "unit": "batch EM iterations/s (64 starts each)", "ms_per_step": loop_ms / K,
            "scaling": "strong",
            "config": {"workload": workload_label("c5", N, p, q, r),
                       "sharding": f"row-sharded {world} x {n_local}, all {C5_STARTS} starts on every rank" + ("" if world == 1 else ", one NCCL allreduce of the starts' moment vectors per iteration"),
                       "path": f"all {C5_STARTS * r} components per data pass (two FP64 GEMMs, em_batch.cu)",
                       "wall_ms_per_step_incl_setup": wall_ms / K},
            "clocks": clocks, "gpu_launches": int(tm["kernel_launches"]),
            "roofline": {"bound": "fp64 (tensor DMMA / FMA pipe; SURVEY.md 8d config 5)", "achieved": flops / (loop_ms * 1e-3) / 1e12 / world,
                         "peak": fp64_peak, "unit": "TFLOP/s per GPU", "frac": flops / (loop_ms * 1e-3) / 1e12 / world / fp64_peak, "traffic": None,
                         "peak_source": "measured in this run: ppls_probe(FP64_DMMA), back-to-back mma.sync m8n8k4 f64 from registers, 8 warps/SM "
                                        "(nominal 148 SMs x 64 FMA/clk x 1.965 GHz = 37.2)",
                         "frac_of_nominal_37": flops / (loop_ms * 1e-3) / 1e12 / world / 37.2,
                         "hbm_gbs_algorithmic": tm["bytes_per_pass"] * tm["stream_launches"] / (loop_ms * 1e-3) / 1e9},
            "loglik_best": best}


def cpu_reference_time(X, Y, start, steps):
    """Time `steps` EM iterations of the op-for-op numpy restatement of the reference."""
    from oracle import ppls_ref as R
    W0, C0, B0, T0 = R.normalise(start["W"], start["C"], start["B"], start["sigT"])
    t0 = time.perf_counter()
    out = R.ppls_simult_loop(X, Y, W0, C0, B0, start["sigE"], start["sigF"], start["sigH"], T0, EMsteps=steps, atol=-np.inf)
    return (time.perf_counter() - t0) / steps, out


def cpu_rows_for(args, N, p, q):
    """Rows of the bounded CPU sample: --cpu-rows (default 1e6), capped by what the op-for-op
    port's temporaries (about 8 copies of an N x (p+q) array) fit into 60 % of the host memory still available."""
    cap = int(0.6 * host_budget_bytes() / (8 * 8 * (p + q))) if host_budget_bytes() else args.cpu_rows
    return max(min(N, args.cpu_rows, max(cap, 10_000)), 1)


def cpu_baseline(args, name, truth, start, N, p, q, r):
    from oracle import cbind, ppls_fused as F
    L = cbind.load()
    blas, omp = set_cpu_threads(L)
    ns = cpu_rows_for(args, N, p, q)
    X, Y = cbind.gen_data(L, SEED, 0, ns, truth)
    steps = 1 if ns * (p + q) > 2e8 else (2 if ns * (p + q) > 2e7 else 5)
    per_iter, _ = cpu_reference_time(X, Y, start, steps)
    t0 = time.perf_counter()
    F.fused_loop(L, X, Y, start["W"], start["C"], start["B"], start["sigE"], start["sigF"], start["sigH"], start["sigT"],
                 EMsteps=steps, atol=-np.inf, want_mu=False)
    fused_iter = (time.perf_counter() - t0) / (steps + 1)
    scale = N / ns
    return {"value": 1.0 / (per_iter * scale), "unit": UNIT, "cores": blas, "kind": "port",
            "sample": f"oracle/ppls_ref.py (op-for-op numpy restatement of Expect_M/Maximiz_M/logl_W; BLAS threads = {blas} of "
                      f"{host_cores()} host cores) on the first {ns} rows of the same generator, {steps} EM step(s), time scaled "
                      f"x{scale:.0f} to N={N}",
            "seconds_per_step_sample": per_iter,
            "fused_cpu_value": 1.0 / (fused_iter * scale), "fused_cpu_cores": omp,
            "fused_cpu_note": f"oracle/ppls_fused.py: single-pass OpenMP restatement (best-effort CPU, {omp} threads), same sample and scaling"}


def e2e_single_process(env, name, K, args):
    """N > 1, rank 0 only (the other ranks have released their data and wait): the SAME end-to-end call made the way ONE
    R session makes it -- a single process driving all N GPUs through ppls_ctx_create_multi (Context.multi): pinned host X, Y
    -> rows split over the devices and uploaded concurrently -> K EM steps with the peer exchange -> estimates back."""
    import torch
    from ppls_b200 import Context, Data, Fit, synth
    N, p, q, r = WORKLOADS[name]
    if args.rows and name == args.workload:
        N = args.rows
    world = env.world
    truth = synth.truth_params(p, q, r)
    start = synth.start_params(truth, seed=1)
    full = args.e2e_rows == 0 and 8 * N * (p + q) <= 0.55 * host_budget_bytes()
    ns = N if full else min(N, (args.e2e_rows or 1_000_000) * world)
    mctx = Context.multi(list(range(world)))
    pinned = None
    try:
        try:
            pinned = PinnedPair(ns, p, q)
        except Exception as ex:
            print(f"[bench] single-process e2e: pinned buffers unavailable ({ex})", file=sys.stderr)
            return None
        d0 = Data.synthetic(mctx, ns, p, q, truth, seed=SEED)
        d0.export(0, ns, out=(pinned.X, pinned.Y))
        d0.close()
        mctx.sync(); torch.cuda.synchronize()
        e0 = time.perf_counter()
        d2 = Data.from_host(mctx, pinned.X, pinned.Y)
        up_s = time.perf_counter() - e0
        f2 = Fit(d2, start, r, EMsteps=K, atol=-float("inf"), type="SVD")
        f2.run()
        res = f2.result(want_mu=False)
        mctx.sync()
        e_s = time.perf_counter() - e0
        f2.close(); d2.close()
    finally:
        if pinned is not None:
            pinned.close()
        mctx.close()
    scale = N / ns
    return {"value": K / (e_s * scale), "unit": UNIT, "h2d_bytes_per_step": int(8 * N * (p + q) / K),
            "d2h_bytes_per_step": int(8 * ((p + q) * r + 2 * r + 3 + K) / K),
            "path": f"ONE process, {world} GPUs (ppls_ctx_create_multi): upload " + ("the full data" if ns == N else f"{ns} rows, time scaled x{scale:.1f}")
                    + f" from pinned host memory split over the devices, {K} EM steps, download estimates+loglik",
            "full_size": ns == N, "seconds_sample": e_s, "upload_seconds_sample": up_s,
            "upload_gbs_total": 8 * ns * (p + q) / up_s / 1e9, "loglik_last": float(res["loglik"][-1])}


def run_ours(args):
    from ppls_b200 import synth
    env = Env(args)
    K, W = args.steps, max(args.warmup, 0)
    parity = parity_block(env) if env.world > 1 and not args.no_parity else None

    if args.workload == "c5":
        main = bench_c5(env, K, W, args)
    else:
        main = bench_stream(env, args.workload, K, W, args, want_e2e=not args.no_e2e)
    others = {}
    if not args.no_others:
        # at N > 1 C4 runs column-sharded (BASELINE.json's configuration) AND, for comparison, row-sharded (250 rows per GPU at
        # N = 8 through the single-pass wide kernel, one allreduce of the (p+q) x r moment vector per iteration)
        for name in ("c2", "c4", "c5") + (("c4_rows",) if env.world > 1 else ()):
            if name == args.workload:
                continue
            try:
                o = bench_c5(env, K, W, args) if name == "c5" else bench_stream(env, name, K, W, args, want_e2e=False)
                o.pop("e2e", None)
                others[name] = {"it_per_s": o["value"], "unit": o["unit"], "ms_per_step": o["ms_per_step"],
                                "kernel_ms": o["roofline"].get("kernel_ms"), "tail_ms_per_step": o["config"].get("tail_ms_per_step"),
                                "achieved": o["roofline"]["achieved"], "roofline_unit": o["roofline"]["unit"], "peak": o["roofline"]["peak"],
                                "frac": o["roofline"]["frac"], "gpu_launches": o["gpu_launches"], "clocks": o["clocks"],
                                "sharding": o["config"]["sharding"], "workload": o["config"]["workload"],
                                "per_rank_ms_step_k1_tail": o["config"].get("per_rank_ms_step_k1_tail"),
                                "exchange": o["config"].get("exchange")}
            except BaseException as ex:                 # a side workload must not take the headline line down
                others[name] = {"error": f"{type(ex).__name__}: {ex}"}
                env.barrier()
    e2e_sp = None
    if env.world > 1 and not args.no_e2e and args.workload != "c5":
        env.barrier()                                   # every rank has released its data
        if env.rank == 0:
            try:
                e2e_sp = e2e_single_process(env, args.workload, K, args)
            except BaseException as ex:
                e2e_sp = {"error": f"{type(ex).__name__}: {ex}"}
        env.dist.barrier(group=env.host_group)          # the others wait on the host: rank 0's kernels have their GPUs to themselves
    cpu = None
    if env.rank == 0 and not args.no_cpu and env.world == 1:
        N, p, q, r = WORKLOADS[args.workload]
        truth = synth.truth_params(p, q, r)
        cpu = cpu_baseline(args, args.workload, truth, synth.start_params(truth, seed=1), args.rows or N, p, q, r)
    if env.rank == 0:
        out = {"metric": METRIC, "value": main["value"], "unit": main["unit"], "n_gpus": env.world, "steps": K, "warmup": W,
               "ms_per_step": main["ms_per_step"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
               "dtype": "f64", "data": "synthetic", "config": main["config"], "clocks": main["clocks"],
               "gpu_launches": main["gpu_launches"], "roofline": main["roofline"]}
        if main.get("e2e"):
            out["e2e"] = main["e2e"]
        if e2e_sp:
            out["e2e_single_process"] = e2e_sp
        if "loglik_best" in main:
            out["loglik_best"] = main["loglik_best"]
        if cpu:
            out["cpu_baseline"] = cpu
        if parity:
            out["parity"] = parity
        if others:
            out["other_workloads"] = others
        emit(out)
    env.close()


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path on the host cores.  R is not installed, so
    this is the oracle port (oracle/ppls_ref.py: numpy -> multithreaded BLAS, as R's %*% -> BLAS), on a bounded row
    sample of the SAME workload label as the other arm.  Rank 0 alone works; the others exit."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import cbind
    from ppls_b200 import synth
    name = args.workload
    N, p, q, r = WORKLOADS[name]
    if args.rows:
        N = args.rows
    truth = synth.truth_params(p, q, r)
    start = synth.start_params(truth, seed=1)
    L = cbind.load()
    blas, omp = set_cpu_threads(L)
    ns = cpu_rows_for(args, N, p, q)
    X, Y = cbind.gen_data(L, SEED + (4 if name == "c5" else 0), 0, ns, truth)
    K, W = args.steps, max(args.warmup, 0)
    big = ns * (p + q) > 2e7
    K_eff = min(K, 3) if big else K
    if W:
        cpu_reference_time(X, Y, start, 1)
    per_iter, _ = cpu_reference_time(X, Y, start, K_eff)
    scale = N / ns * (C5_STARTS if name == "c5" else 1)      # c5: the reference would run the 64 starts one after another
    value = 1.0 / (per_iter * scale)
    sample = (f"oracle/ppls_ref.py on the first {ns} rows of the same generator, {K_eff} timed EM steps after {min(W, 1)} warm-up "
              f"(bounded from --steps {K} --warmup {W}), BLAS threads = {blas} of {host_cores()} host cores, time scaled x{scale:.0f} to the workload")
    out = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT if name != "c5" else "batch EM iterations/s (64 starts each)",
           "n_gpus": args.gpus, "steps": K, "warmup": W,
           "ms_per_step": per_iter * scale * 1e3, "higher_is_better": True, "scaling": "strong",
           "vs_baseline": None, "dtype": "f64", "data": "synthetic",
           "config": {"workload": workload_label(name, N, p, q, r)},
           "cpu_baseline": {"value": value, "unit": UNIT, "cores": blas, "kind": "port", "sample": sample,
                            "seconds_per_step_sample": per_iter},
           "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(out)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS))
    ap.add_argument("--rows", type=int, default=0, help="override N of the headline workload (debugging)")
    ap.add_argument("--e2e-rows", type=int, default=0, help="rows per GPU of the e2e call; 0 = the full shard when the host can pin it, else 1e6")
    ap.add_argument("--cpu-rows", type=int, default=1_000_000, help="rows of the bounded CPU sample (BASELINE.md section 4)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-others", action="store_true", help="skip the other named shapes (other_workloads)")
    ap.add_argument("--no-parity", action="store_true", help="skip the untimed N > 1 parity block")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
