#!/usr/bin/env python
"""Benchmark of the return / replay hot path on B200 (contract: see the task statement).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

Headline metric (BASELINE.json): GAE transitions/s on configs[1], the synthetic GAE sweep
4096 env streams x 2048 steps per GPU, gamma 0.99, lambda 0.95 (weak scaling: every rank owns
its own 4096 streams, as VectorReplayBuffer sub-buffers shard with no communication).
A "step" is one pass of ``ts_gae_rows`` (C-ABI) over the rank's rollout, resident in HBM.
The same JSON line also carries: ``e2e`` (the drop-in call
``BasePolicy.compute_episodic_return(batch, buffer, indices, v_s_, v_s)`` with NumPy arrays in
and float64 NumPy arrays out, as the reference's callers use it -- host-to-device and
device-to-host copies inside the timed region; ``e2e.pipelined`` is the package's own
pinned-tensor pipeline with f32 outputs and cross-step overlap), ``roofline`` (algorithmic
21 B/transition over the CUDA-event kernel time vs MEASURED_PEAKS.json), ``cpu_baseline`` (the
oracle on the box's host cores, bounded sample: the C port and the numba port), ``per`` (PER
samples/s on configs[3] / configs[4], a complete object of its own: value, roofline,
cpu_baseline, e2e, and at N > 1 a bit-exactness check against the oracle's single global tree),
``updates`` (C3 / C4 through the policy classes; at N > 1 the data-parallel learner) and
``clocks``.

``--impl reference`` times the CPU restatement of the reference (oracle/, the numba kernels
re-stated in C) over the FULL workload of this arm's config on all host cores (explicit thread
count = os.cpu_count(), identical at every N: torchrun's OMP_NUM_THREADS=1 is not inherited); it
is the one place besides ``cpu_baseline`` where bench.py executes oracle/ code, and only as the
thing being compared against.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

E_PER_GPU, T_STEPS = 4096, 2048
GAMMA, LAMBDA = 0.99, 0.95
ALG_BYTES_PER_TRANSITION = 21  # rew 4 + v_s 4 + v_s_ 4 + end flag 1 + adv 4 + ret 4
MOVED_BYTES_PER_TRANSITION = 22  # terminated and truncated are two separate bytes
E2E_CHUNKS = int(os.environ.get("TS_E2E_CHUNKS", "2"))
NCU_DRAM_BYTES_PER_LAUNCH = 145051904  # profiles/r02_gae_ncu.md: 117.49 MB read + 27.57 MB write
PER_LEAVES, PER_BATCH, PER_ALPHA, PER_BETA = 1 << 20, 4096, 0.6, 0.4
PER_ALG_BYTES_PER_SAMPLE = 664


def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=2000)
    p.add_argument("--warmup", type=int, default=20)
    p.add_argument("--impl", default="b200", choices=["b200", "reference"])
    p.add_argument("--streams", type=int, default=E_PER_GPU, help="env streams per GPU")
    p.add_argument("--horizon", type=int, default=T_STEPS)
    p.add_argument("--skip-per", action="store_true")
    p.add_argument("--skip-e2e", action="store_true")
    p.add_argument("--skip-cpu", action="store_true")
    p.add_argument("--skip-updates", action="store_true",
                   help="skip the C3 PPOPolicy.learn / C4 DQNPolicy.update legs")
    return p.parse_args()


def rollout(E, T, seed):
    """SURVEY.md 8d C2 generator (same as tests/helpers.c2_rollout)."""
    rng = np.random.default_rng(seed)
    rew = rng.random((E, T), dtype=np.float32)
    v_s = rng.standard_normal((E, T), dtype=np.float32)
    v_next = np.empty_like(v_s)
    v_next[:, :-1] = v_s[:, 1:]
    v_next[:, -1] = rng.standard_normal(E, dtype=np.float32)
    term = rng.random((E, T)) < 1 / 200
    trunc = (rng.random((E, T)) < 1 / 1000) & ~term
    return rew, v_s, v_next, term.view(np.uint8), trunc.view(np.uint8)


def note(msg: str) -> None:
    """Progress on stderr (rank-tagged): tells where a multi-rank run stopped."""
    print(f"[bench r{os.environ.get('RANK', '0')} {time.strftime('%H:%M:%S')}] {msg}",
          file=sys.stderr, flush=True)


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def host_info():
    """CPU model, core count and the versions that pin a CPU baseline."""
    model = "unknown"
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("model name"):
                    model = line.split(":", 1)[1].strip()
                    break
    except OSError:
        pass
    info = {"cpu_model": model, "host_cpus": os.cpu_count(), "numpy": np.__version__}
    try:
        import numba
        info["numba"] = numba.__version__
    except Exception:
        info["numba"] = None
    return info


class ClockSampler:
    """Samples SM clock and throttle reasons through NVML while a timed region runs."""

    REASONS = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown",
               0x40: "hw_thermal_slowdown", 0x80: "hw_power_brake_slowdown"}

    def __init__(self, index: int) -> None:
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thread = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:  # NVML unavailable: clocks stay unknown
            self.nv = None

    def _loop(self):
        nv = self.nv
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                mask = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                for bit, name in self.REASONS.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.002)

    def __enter__(self):
        if self.nv is not None:
            self._stop.clear()
            self._thread = threading.Thread(target=self._loop, daemon=True)
            self._thread.start()
        return self

    def __exit__(self, *exc):
        if self._thread is not None:
            self._stop.set()
            self._thread.join()
            self._thread = None

    def summary(self):
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(s)}

    def bad(self):
        return bool(self.reasons & {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"})


# ---------------------------------------------------------------------------------------------
# CPU side: the oracle port of the reference (C restatement of the numba kernels)
# ---------------------------------------------------------------------------------------------
def cpu_gae(threads: int, seconds: float, rows: int, T: int):
    """Times orc_episodic_return_rows on a bounded sample (rows x T); returns transitions/s."""
    from oracle import oracle as orc
    rew, v_s, v_next, term, trunc = rollout(rows, T, seed=1234)
    orc.episodic_return_rows(rew, v_s, v_next, term, trunc, GAMMA, LAMBDA, threads)  # warm
    n, t0 = 0, time.perf_counter()
    times = []
    while time.perf_counter() - t0 < seconds or n < 3:
        a = time.perf_counter()
        orc.episodic_return_rows(rew, v_s, v_next, term, trunc, GAMMA, LAMBDA, threads)
        times.append(time.perf_counter() - a)
        n += 1
    med = float(np.median(times))
    return rows * T / med, med, n


def cpu_per(seconds: float):
    """Times one PER step (sample + weights + update) of the oracle port; samples/s."""
    from oracle import oracle as orc
    rng = np.random.default_rng(2)
    tree = orc.SegTree(PER_LEAVES)
    tree[np.arange(PER_LEAVES)] = (np.abs(rng.standard_normal(PER_LEAVES)) + 1e-7) ** PER_ALPHA
    mx = mn = 1.0
    times, t0 = [], time.perf_counter()
    while time.perf_counter() - t0 < seconds or len(times) < 5:
        u = rng.random(PER_BATCH)
        td = rng.standard_normal(PER_BATCH).astype(np.float32)
        a = time.perf_counter()
        _, _, mx, mn = orc.per_step_c(tree, u, td, PER_ALPHA, PER_BETA, True, mx, mn)
        times.append(time.perf_counter() - a)
    med = float(np.median(times))
    return PER_BATCH / med, med, len(times)


def cpu_gae_numba(seconds: float, rows: int, T: int):
    """The reference's own technology on one core: numba _gae_return + the NumPy glue of
    compute_episodic_return (oracle/oracle_nb.py, calibrated against the imported reference in
    profiles/r02_reference_cpu.json: 1.01x its time at C2).  transitions/s, or None without numba."""
    try:
        from oracle import oracle_nb as nb
        nb.warm()
    except Exception:
        return None
    rew, v_s, v_next, term, trunc = rollout(rows, T, seed=1234)
    n = rows * T
    r64 = rew.reshape(-1).astype(np.float64)  # the reference stores rewards as f64 (manager.py:150)
    t, c = term.reshape(-1).astype(bool), trunc.reshape(-1).astype(bool)
    idx = np.arange(n)
    unf = np.flatnonzero(~(t | c).reshape(rows, T)[:, -1]) * T + T - 1
    vs, vn = v_s.reshape(-1), v_next.reshape(-1)
    f = lambda: nb.compute_episodic_return(r64, t, c, t, idx, unf, vn, vs, GAMMA, LAMBDA)  # noqa: E731
    f()
    times, t0 = [], time.perf_counter()
    while time.perf_counter() - t0 < seconds or len(times) < 3:
        a = time.perf_counter()
        f()
        times.append(time.perf_counter() - a)
    med = float(np.median(times))
    return n / med, med, len(times)


def cpu_per_numba(seconds: float):
    """One PER step the way the reference runs it (prio.py:59-102 over segtree.py:87-137: numba
    array kernels + NumPy), one core; samples/s or None without numba."""
    try:
        from oracle import oracle_nb as nb
        nb.warm()
    except Exception:
        return None
    rng = np.random.default_rng(2)
    per = nb.PER(PER_LEAVES, PER_ALPHA, PER_BETA)
    per.set_leaves(np.arange(PER_LEAVES),
                   (np.abs(rng.standard_normal(PER_LEAVES)) + 1e-7) ** PER_ALPHA)
    tds = [rng.standard_normal(PER_BATCH).astype(np.float32) for _ in range(8)]
    times, t0, i = [], time.perf_counter(), 0
    while time.perf_counter() - t0 < seconds or len(times) < 5:
        a = time.perf_counter()
        idx = per.sample_indices(PER_BATCH)
        per.batch_weight(idx)
        per.update_weight(idx, tds[i % 8])
        times.append(time.perf_counter() - a)
        i += 1
    med = float(np.median(times))
    return PER_BATCH / med, med, len(times)


def run_reference(args):
    """--impl reference: the reference's CPU algorithm on the host cores (rank 0 only; the other
    ranks exit without work).  Each step is one pass over the FULL workload of this arm's config
    (args.streams x args.horizon transitions); the thread count is explicit and the same at
    every N."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    from oracle import oracle as orc
    threads = max(1, os.cpu_count() or 1)
    rows = args.streams
    rew, v_s, v_next, term, trunc = rollout(rows, args.horizon, seed=1234)
    f = lambda th: orc.episodic_return_rows(rew, v_s, v_next, term, trunc, GAMMA, LAMBDA, th)  # noqa: E731
    for _ in range(max(args.warmup, 1)):
        f(threads)
    steps = max(1, min(args.steps, 200))
    t0 = time.perf_counter()
    for _ in range(steps):
        f(threads)
    dt = time.perf_counter() - t0
    value = rows * args.horizon * steps / dt
    one, _, _ = cpu_gae(1, 2.0, min(rows, 256), args.horizon)
    nb1 = cpu_gae_numba(3.0, min(rows, 512), args.horizon)
    per, per_med, _ = cpu_per(3.0)
    per_nb = cpu_per_numba(3.0)
    info = host_info()
    line = {
        "impl": "reference", "metric": "GAE transitions/s", "value": value,
        "unit": "transitions/s", "n_gpus": args.gpus, "steps": steps, "warmup": args.warmup,
        "ms_per_step": dt / steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64 arithmetic, f32 storage", "data": "synthetic",
        "config": {"workload": "GAE sweep 4096 streams x 2048 steps per GPU, gamma 0.99 "
                               "lambda 0.95 (BASELINE.json configs[1])",
                   "streams_per_gpu": args.streams, "horizon": args.horizon},
        "cpu_baseline": {"value": value, "unit": "transitions/s", "cores": threads,
                         "kind": "port",
                         "sample": f"the full workload: {rows} streams x {args.horizon} steps per "
                                   f"pass, OpenMP over streams, {threads} threads (explicit, not "
                                   f"OMP_NUM_THREADS); C port 1 thread: {one:.3e}; numba port "
                                   "(the reference's technology) 1 core: "
                                   + (f"{nb1[0]:.3e}" if nb1 else "numba unavailable"),
                         "port_1thread": one, "numba_port_1core": nb1[0] if nb1 else None,
                         **info},
        "e2e": {"value": value, "unit": "transitions/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "per": {"metric": "PER samples/s", "value": per, "unit": "samples/s", "cores": 1,
                "us_per_step": per_med * 1e6, "kind": "port",
                "numba_port": ({"value": per_nb[0], "us_per_step": per_nb[1] * 1e6}
                               if per_nb else None),
                "workload": "2^20 leaves, batch 4096, sample + weights + update_weight"},
        "gpu_launches": 0,
    }
    emit_result(line)


# ---------------------------------------------------------------------------------------------
# GPU side
# ---------------------------------------------------------------------------------------------
def run_b200(args):
    import torch
    import torch.distributed as dist

    from tianshou_fork_b200 import _lib as L

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device(f"cuda:{local}")
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lib = L.lib()  # raises if the CUDA extension is missing: there is no fallback
    E, T = args.streams, args.horizon
    K, W = args.steps, max(args.warmup, 3)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- inputs: pinned host copies (for e2e) and HBM-resident copies (for value) -------------
    host = [torch.from_numpy(a).pin_memory() for a in rollout(E, T, seed=rank)]
    d_rew, d_vs, d_vn, d_term, d_trunc = [h.to(dev) for h in host]
    adv = torch.empty(E, T, dtype=torch.float32, device=dev)
    ret = torch.empty(E, T, dtype=torch.float32, device=dev)
    # Successive launches work on DISTINCT rollouts and outputs (ROTATE sets, 185 MB each, well
    # over the 126 MB L2 together), so that nothing a launch reads or writes can still be in L2
    # from the launch before; re-launching on one set measures ~3 % faster.
    rotate = max(1, int(os.environ.get("TS_BENCH_ROTATE", "3")))
    sets = [((d_rew, d_vs, d_vn, d_term, d_trunc), adv, ret)]
    for r in range(1, rotate):
        ins = tuple(torch.from_numpy(a).to(dev) for a in rollout(E, T, seed=1000 * r + rank))
        sets.append((ins, torch.empty_like(adv), torch.empty_like(ret)))
    stream = torch.cuda.current_stream()
    launches = 0
    step_no = [0]

    def gae_step(s=None):
        ins, o_adv, o_ret = sets[step_no[0] % rotate]
        step_no[0] += 1
        L.check(lib.ts_gae_rows(
            ins[0].data_ptr(), ins[1].data_ptr(), ins[2].data_ptr(), ins[3].data_ptr(),
            ins[4].data_ptr(), E, T, T, T, GAMMA, LAMBDA, L.TS_F32, L.TS_F32, L.TS_F32,
            o_adv.data_ptr(), o_ret.data_ptr(), s if s is not None else stream.cuda_stream),
            "ts_gae_rows")

    sampler = ClockSampler(local)
    # The K timed steps are K launches of the same C-ABI call; they are issued as replays of a
    # CUDA graph of GRAPH_STEPS launches (plus an eager remainder) so that the host never sits
    # between two launches.  TS_BENCH_GRAPH_STEPS=1 times the plain eager loop instead.
    graph_steps = max(1, min(K, int(os.environ.get("TS_BENCH_GRAPH_STEPS", "100"))))
    graph = None
    if graph_steps > 1:
        try:
            side = torch.cuda.Stream(dev)
            side.wait_stream(stream)
            with torch.cuda.stream(side):
                for _ in range(3):
                    gae_step(side.cuda_stream)
            side.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=side):
                for _ in range(graph_steps):
                    gae_step(side.cuda_stream)
        except Exception as e:  # pragma: no cover - capture is expected to work on the GPU box
            print(f"graph capture of the GAE loop failed ({e}); timing the eager loop",
                  file=sys.stderr)
            graph, graph_steps = None, 1

    def run_steps(n):
        if graph is not None:
            for _ in range(n // graph_steps):
                graph.replay()
            n = n % graph_steps
        for _ in range(n):
            gae_step()

    def timed_gae():
        run_steps(max(W, graph_steps if graph is not None else W))
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with sampler:
            e0.record(stream)
            run_steps(K)
            e1.record(stream)
            barrier()
        return e0.elapsed_time(e1)  # ms on the launching stream

    note("gae device-resident leg")
    ms = timed_gae()
    if sampler.bad():  # thermal / hw slowdown seen: re-measure once
        sampler.samples.clear()
        sampler.reasons.clear()
        ms = timed_gae()
    launches += K
    ms = max_over_ranks(ms)
    n_trans = E * T
    value = n_trans * world * K / (ms * 1e-3)
    kernel_s = ms * 1e-3 / K
    peak, peak_src = measured_peaks()
    achieved = n_trans * ALG_BYTES_PER_TRANSITION / kernel_s / 1e9
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": NCU_DRAM_BYTES_PER_LAUNCH if (E, T) == (
                    E_PER_GPU, T_STEPS) else None,
                "traffic_source": "ncu --set full, profiles/r02_gae_ncu.md (r02_gae_top): dram read "
                                  "117.49 MB + write 27.57 MB; the rest of the 67.1 MB output is "
                                  "still dirty in L2 at kernel end",
                "peak_source": peak_src,
                "kernel": "gae_rows_tma_kernel<float,float,float,256> (persistent, cp.async.bulk pipelined, 3 stages x 2 CTAs/SM)",
                "alg_bytes_per_launch": n_trans * ALG_BYTES_PER_TRANSITION,
                "moved_bytes_per_launch": n_trans * MOVED_BYTES_PER_TRANSITION,
                "us_per_launch": kernel_s * 1e6}

    # ---- e2e: the drop-in call with host arrays in and out ------------------------------------------
    e2e = None
    if not args.skip_e2e:
        note("e2e leg")
        e2e = run_e2e(torch, dist, L, args, dev, rank, world, host, sets, barrier, max_over_ranks,
                      sampler, K)

    # ---- PER: sample + importance weights + update_weight, CUDA-graph replay ------------------
    per = None
    if not args.skip_per:
        note("per leg")
        per = run_per(torch, dist, L, lib, dev, rank, world, barrier, max_over_ranks)

    # ---- CPU baseline (rank 0, N == 1 only): the oracle on the box's host cores -----------------
    cpu = None
    if rank == 0 and world == 1 and not args.skip_cpu:
        v1, med1, n1 = cpu_gae(1, 6.0, 512, T)
        nbr = cpu_gae_numba(6.0, 512, T)
        cpu = {"value": v1, "unit": "transitions/s", "cores": 1, "kind": "port",
               "sample": f"512 streams x {T} steps per pass, {n1} passes, median "
                         f"{med1 * 1e3:.2f} ms; oracle/ts_oracle.c (C restatement of numba "
                         "_gae_return + value mask + end flags), f32 in / f64 out",
               "numba_port": ({"value": nbr[0], "unit": "transitions/s", "cores": 1,
                               "sample": f"512 streams x {T} steps per pass, {nbr[2]} passes, "
                                         f"median {nbr[1] * 1e3:.2f} ms; oracle/oracle_nb.py: numba "
                                         "_gae_return + the NumPy glue of compute_episodic_return "
                                         "(np.isin, value mask, f64 temporaries) as the reference "
                                         "runs it; 1.01x the imported reference's time in the "
                                         "build container (profiles/r02_reference_cpu.json)"}
                              if nbr else None),
               **host_info()}
        if per is not None:
            pv, pmed, pn = cpu_per(3.0)
            per["cpu_baseline"] = {"value": pv, "unit": "samples/s", "cores": 1, "kind": "port",
                                   "sample": f"{pn} steps, median {pmed * 1e6:.1f} us "
                                             "(oracle/ts_oracle.c orc_per_step)", **host_info()}
            per["speedup_vs_cpu_port"] = per["value"] / pv
            pnb = cpu_per_numba(3.0)
            if pnb:
                per["cpu_baseline"]["numba_port"] = {
                    "value": pnb[0], "unit": "samples/s", "cores": 1,
                    "sample": f"{pnb[2]} steps, median {pnb[1] * 1e6:.1f} us; oracle/oracle_nb.py: "
                              "the reference's numba array kernels + NumPy (0.64x the imported "
                              "reference's step time in the build container, i.e. faster than it)"}
                per["speedup_vs_numba_port"] = per["value"] / pnb[0]

    # ---- C3 / C4 through the public API (rank 0, N == 1 only; reported beside the headline) ----
    updates = None
    if rank == 0 and world == 1 and not args.skip_updates and not args.skip_e2e:
        updates = run_updates()
    elif world > 1 and not args.skip_updates:
        # the data-parallel learner (all ranks take part: one NCCL all-reduce per minibatch)
        sys.path.insert(0, os.path.join(ROOT, "scripts"))
        import bench_dp
        note("data-parallel updates leg")
        updates = bench_dp.run_dp_updates(torch, dist, dev, rank, world, max_over_ranks)

    if rank == 0:
        line = {
            "metric": "GAE transitions/s", "value": value, "unit": "transitions/s",
            "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms / K,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64 arithmetic, f32 storage", "data": "synthetic",
            "config": {"workload": "GAE sweep 4096 streams x 2048 steps per GPU, gamma 0.99 "
                                   "lambda 0.95 (BASELINE.json configs[1])",
                       "streams_per_gpu": E, "horizon": T, "storage": "f32 values, u8 flags",
                       "l2": f"inputs+outputs {n_trans * 22 / 1e6:.0f} MB per step > 126 MB L2, and "
                             f"successive steps rotate over {rotate} distinct input/output sets "
                             f"({rotate * n_trans * 22 / 1e6:.0f} MB): no reuse across steps, no "
                             "flush needed",
                       "timed_loop": (f"CUDA graph of {graph_steps} launches of ts_gae_rows, replayed"
                                      if graph is not None else "eager launches of ts_gae_rows"),
                       "parallelism": f"env streams sharded over {world} GPU(s), no collective"},
            "clocks": sampler.summary(), "e2e": e2e, "gpu_launches": launches,
            "roofline": roofline, "cpu_baseline": cpu, "per": per, "updates": updates,
        }
        emit_result(line)
    if world > 1:
        # Every leg is done and the line is out.  Tearing NCCL down with captured graphs that
        # hold collectives and symmetric-memory mappings still alive blocked for minutes in
        # destroy_process_group() (r02o: all results printed, then a 20-minute wait for the
        # watchdog); the ranks leave together instead.
        note("done")
        barrier()
        sys.stderr.flush()
        os._exit(0)


def run_e2e(torch, dist, L, args, dev, rank, world, host, sets, barrier, max_over_ranks, sampler,
            K):
    """End to end through the reference-facing call.

    The rollout sits in a device-resident ``VectorReplayBuffer`` (filled through ``add``);
    ``indices`` is what ``buffer.sample_indices(0)`` returns.  Every timed step calls
    ``BasePolicy.compute_episodic_return(batch, buffer, indices, v_s_, v_s, gamma, gae_lambda)``
    (tianshou/policy/base.py:337-346) with ``batch.rew / terminated / truncated`` and the critic
    values as NumPy arrays (views of pinned host memory, so the copies are DMA from pinned
    memory as the contract asks) and gets ``(returns, advantage)`` back as float64 NumPy arrays
    like numba's.  Host-to-device and device-to-host copies are inside the timed region; the call
    is synchronous (it returns host arrays), so successive steps do not overlap."""
    import tianshou_fork_b200 as ts
    from tianshou_fork_b200.pipeline import GaeHostPipeline, gae_host_call
    E, T = args.streams, args.horizon
    n_trans = E * T
    stream = torch.cuda.current_stream()
    rew_h, vs_h, vn_h, term_h, trunc_h = host  # pinned torch tensors [E, T]
    buf = ts.VectorReplayBuffer(E * T, E, rew_dtype=torch.float32)
    z = np.zeros(E, np.uint8)
    rew_np, term_np, trunc_np = rew_h.numpy(), term_h.numpy().view(bool), trunc_h.numpy().view(bool)
    for t in range(T):
        buf.add(ts.Batch(obs=z, act=z, rew=rew_np[:, t], terminated=term_np[:, t],
                         truncated=trunc_np[:, t], obs_next=z, info={}))
    indices = buf.sample_indices(0)
    assert buf.is_canonical(indices) == (E, T)
    batch_np = ts.Batch(rew=rew_np.reshape(-1), terminated=term_np.reshape(-1),
                        truncated=trunc_np.reshape(-1))
    vs_np, vn_np = vs_h.numpy().reshape(-1), vn_h.numpy().reshape(-1)
    hc = gae_host_call(dev)

    def call(batch=batch_np, vn=vn_np, vs=vs_np, **kw):
        return ts.BasePolicy.compute_episodic_return(batch, buf, indices, vn, vs, GAMMA, LAMBDA,
                                                     **kw)

    Ke = max(3, min(K, 50))
    for _ in range(3):
        ret, adv = call()
    assert isinstance(adv, np.ndarray) and adv.dtype == np.float64
    if rank == 0:  # the copied-back result is the real one (f64 here, f32 from the device leg)
        d_adv = sets[0][1].cpu().numpy().reshape(-1)
        assert np.allclose(adv, d_adv, rtol=1e-5, atol=1e-4)
    barrier()
    h0, d0, l0 = hc.h2d_bytes, hc.d2h_bytes, hc.launches
    with sampler:
        t0 = time.perf_counter()
        for _ in range(Ke):
            ret, adv = call()
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        barrier()
    ems = max_over_ranks(dt * 1e3)
    h2d, d2h = (hc.h2d_bytes - h0) // Ke, (hc.d2h_bytes - d0) // Ke
    launches = hc.launches - l0

    def wall(fn, n=5):
        fn()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(n):
            fn()
        torch.cuda.synchronize()
        return max_over_ranks((time.perf_counter() - t0) / n * 1e3)

    # variants of the same call (context, not the headline)
    ms_f32 = wall(lambda: call(out_dtype=torch.float32))
    dev_batch = ts.Batch(rew=buf.rew, terminated=buf.terminated, truncated=buf.truncated)
    ms_devbatch = wall(lambda: call(batch=dev_batch))
    pageable = [np.array(a) for a in (vn_np, vs_np)]
    pb = ts.Batch(rew=np.array(batch_np.rew), terminated=np.array(batch_np.terminated),
                  truncated=np.array(batch_np.truncated))
    ms_pageable = wall(lambda: call(batch=pb, vn=pageable[0], vs=pageable[1]), n=3)

    # ---- the package's own pinned-tensor pipeline (f32 out, successive steps overlap) ----------
    pipe = GaeHostPipeline(E, T, chunks=E2E_CHUNKS, device=dev)
    pipe.overlap = os.environ.get("TS_E2E_OVERLAP", "1") == "1"
    h_adv = torch.empty(E, T, dtype=torch.float32).pin_memory()
    h_ret = torch.empty(E, T, dtype=torch.float32).pin_memory()
    for _ in range(3):
        pipe.run(*host, h_adv, h_ret, GAMMA, LAMBDA)
    pipe.join(stream)
    barrier()
    Kp = max(3, min(K, 50))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(Kp):
        pipe.run(*host, h_adv, h_ret, GAMMA, LAMBDA)
    pipe.join(stream)  # every step's device->host copy has landed before the stop event
    e1.record(stream)
    barrier()
    pms = max_over_ranks(e0.elapsed_time(e1))

    # ---- what the link can do: one big pinned copy each way, both at once; all ranks at once -----
    d_rew, d_adv_t = sets[0][0][0], sets[0][1]

    def link_gbs(fn, nbytes):
        best = 1e9
        for _ in range(3):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            barrier()  # N > 1: every rank copies at the same time (shared PCIe switches / NUMA)
            a.record(stream)
            fn()
            b.record(stream)
            torch.cuda.synchronize()
            best = min(best, a.elapsed_time(b))
        return nbytes / (best * 1e-3) / 1e9

    s2 = torch.cuda.Stream(dev)
    nb_in, nb_out = rew_h.numel() * 4, h_adv.numel() * 4

    def both():
        s2.wait_stream(stream)
        with torch.cuda.stream(s2):
            h_adv.copy_(d_adv_t, non_blocking=True)
        d_rew.copy_(rew_h, non_blocking=True)
        stream.wait_stream(s2)

    link = {"h2d_GBps": link_gbs(lambda: d_rew.copy_(rew_h, non_blocking=True), nb_in),
            "d2h_GBps": link_gbs(lambda: h_adv.copy_(d_adv_t, non_blocking=True), nb_out),
            "duplex_GBps": link_gbs(both, nb_in + nb_out),
            "how": "134 MB pinned copies, best of 3, per rank with all ranks copying concurrently"}

    def total(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t)
        return float(t.item())

    for k in ("h2d_GBps", "d2h_GBps", "duplex_GBps"):
        link["aggregate_" + k] = total(link[k])
    link["floor_ms_per_step"] = max(h2d / link["h2d_GBps"], d2h / link["d2h_GBps"],
                                    (h2d + d2h) / link["duplex_GBps"]) / 1e6
    return {
        "value": n_trans * world * Ke / (ems * 1e-3), "unit": "transitions/s",
        "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "ms_per_step": ems / Ke,
        "steps": Ke, "launches_per_step": launches // Ke,
        "api": "BasePolicy.compute_episodic_return(batch, buffer, indices, v_s_, v_s, gamma, "
               "gae_lambda): batch.rew f32 / terminated / truncated and v_s_, v_s f32 as NumPy "
               "arrays (pinned host memory) in, (returns, advantage) float64 NumPy arrays out; "
               "buffer = device-resident VectorReplayBuffer, indices = buffer.sample_indices(0); "
               "synchronous call, wall clock, max over ranks",
        "variants_ms_per_step": {
            "float32_out (out_dtype=torch.float32, this package's extension)": ms_f32,
            "batch fields already on the device (batch = buffer storage), NumPy values": ms_devbatch,
            "pageable NumPy arrays (bounced through pinned buffers by the call)": ms_pageable},
        "pipelined": {
            "value": n_trans * world * Kp / (pms * 1e-3), "unit": "transitions/s",
            "ms_per_step": pms / Kp, "h2d_bytes_per_step": pipe.h2d_bytes,
            "d2h_bytes_per_step": pipe.d2h_bytes,
            "api": f"tianshou_fork_b200.pipeline.GaeHostPipeline.run: pinned torch tensors, f32 "
                   f"outputs, {E2E_CHUNKS} row chunks, successive steps overlap on the link "
                   "(not the reference's signature)"},
        "link": link,
    }


def run_updates():
    """BASELINE.json configs[2] and configs[3] through the public API: wall time per minibatch
    of PPOPolicy.learn (Ant shape) and per DQNPolicy.update (Atari shape, PER 2^20), eager and
    CUDA-graphed, with the reference's learn() arithmetic in torch eager on the same GPU and on
    the host beside it (scripts/bench_rows.py row_learn / row_dqn).  Never fails the bench."""
    try:
        sys.path.insert(0, os.path.join(ROOT, "scripts"))
        import bench_rows
        bench_rows.ROWS.clear()
        bench_rows.row_learn()
        bench_rows.row_dqn()
        out = {"unit": "us (wall, per minibatch / per update)"}
        for r in bench_rows.ROWS:
            if r["row"].startswith("C3"):
                out[f"c3_ppo_learn_mb{r['minibatch']}"] = {
                    "cuda_graph": r["this_cuda_graph_us"], "eager": r["this_eager_us"],
                    "reference_arithmetic_same_gpu": r["reference_arithmetic_torch_gpu_us"],
                    "reference_arithmetic_cpu": r["reference_arithmetic_torch_cpu_us"],
                    "cpu_threads": r["cpu_threads"]}
            else:
                out[f"c4_dqn_update_bsz{r['bsz']}"] = {
                    "cuda_graph": r["update_cuda_graph_wall_us"], "eager": r["update_wall_us"],
                    "networks_only_eager": r["networks_only_us"]}
        return out
    except Exception as e:  # reported, not fatal: the headline legs above are already measured
        return {"error": f"{type(e).__name__}: {e}"}


def run_per(torch, dist, L, lib, dev, rank, world, barrier, max_over_ranks):
    """PER samples/s, the metric's second half: scripts/bench_per.py (strong + weak scaling at
    N > 1, parity against the oracle's single global tree, e2e through the buffer classes)."""
    sys.path.insert(0, os.path.join(ROOT, "scripts"))
    import bench_per
    return bench_per.run_per(torch, dist, L, lib, dev, rank, world, barrier, max_over_ranks,
                             measured_peaks()[0])


_RESULT_FD = None


def _keep_stdout_clean():
    """The driver reads ONE JSON line from stdout.  Libraries (NCCL's version banner, torchrun
    children) write to file descriptor 1 behind Python's back: point fd 1 at stderr for the
    whole run and keep a private duplicate of the real stdout for the result line."""
    global _RESULT_FD
    sys.stdout.flush()
    _RESULT_FD = os.dup(1)
    os.dup2(2, 1)


def emit_result(line: dict) -> None:
    data = (json.dumps(line) + "\n").encode()
    if _RESULT_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_RESULT_FD, data)


def main():
    args = parse_args()
    _keep_stdout_clean()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
