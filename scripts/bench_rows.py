#!/usr/bin/env python
"""Per-row timings of the hot path on the GPU box, each beside the
CPU oracle port on one host core.  Prints one JSON object per row; results are committed under
profiles/.  GPU times are CUDA-event timings through the package's public API (host launch
overhead included) unless the row says "graph".

    python scripts/bench_rows.py > gpurun_out/rows.jsonl
"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import tianshou_fork_b200 as ts  # noqa: E402
from oracle import oracle as orc  # noqa: E402
from tianshou_fork_b200 import _lib as L  # noqa: E402

DEV = torch.device("cuda:0")
PEAK = 6545.3
try:
    PEAK = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass


def gpu_us(fn, iters=200, warm=10):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    best = 1e30
    for _ in range(3):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(iters):
            fn()
        b.record()
        torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b) * 1e3 / iters)
    return best


def graph_us(fn, calls=8, replays=50):
    side = torch.cuda.Stream(DEV)
    with torch.cuda.stream(side):
        for _ in range(3):
            fn(side.cuda_stream)
    side.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=side):
        for _ in range(calls):
            fn(side.cuda_stream)
    return gpu_us(g.replay, iters=replays, warm=5) / calls


def cpu_us(fn, seconds=2.0):
    fn()
    t, n = [], 0
    t0 = time.perf_counter()
    while time.perf_counter() - t0 < seconds or n < 3:
        a = time.perf_counter()
        fn()
        t.append(time.perf_counter() - a)
        n += 1
    return float(np.median(t)) * 1e6


ROWS = []  # every emitted row, for callers that import this module (bench.py)


def emit(row, **kw):
    ROWS.append(dict(row=row, **kw))
    print(json.dumps(ROWS[-1]), flush=True)


def fill_vec(buf, E, T, rng, p_term=1 / 1000, p_trunc=1 / 10000, rew3=True):
    """Bulk-fill a device VectorReplayBuffer through its storage (rollouts are synthetic)."""
    n = E * T
    rew = rng.choice(np.array([-1.0, 0.0, 1.0]), n, p=[0.05, 0.9, 0.05]) if rew3 else rng.random(n)
    term = rng.random(n) < p_term
    trunc = (rng.random(n) < p_trunc) & ~term
    for t in range(2):  # two real adds allocate the storage and initialise bookkeeping
        buf.add(ts.Batch(obs=np.zeros((E, 1), np.float32), act=np.zeros(E, np.int64),
                         rew=np.zeros(E), terminated=np.zeros(E, bool), truncated=np.zeros(E, bool),
                         obs_next=np.zeros((E, 1), np.float32), info={}))
    buf._meta.rew.copy_(torch.from_numpy(rew).to(buf._meta.rew.dtype))
    buf._meta.terminated.copy_(torch.from_numpy(term))
    buf._meta.truncated.copy_(torch.from_numpy(trunc))
    buf._meta.done.copy_(torch.from_numpy(term | trunc))
    buf._rings[:, 0] = 0
    buf._rings[:, 1] = T
    buf._lengths[:] = T
    buf.last_index[:] = np.arange(E) * T + T - 1
    buf._version += 1
    for b in buf.buffers:
        b._index, b._size = 0, T
    return rew, term, trunc


def row_gae_api():
    """a1 through the drop-in API: BasePolicy.compute_episodic_return(batch, buffer, indices,
    v_s_, v_s) on a device VectorReplayBuffer (sample(0), CUDA tensors in and out), beside the
    oracle port on one core and the survey's probe of the reference (2.38 ms at C3)."""
    for E, T in ((64, 2048), (4096, 2048)):
        rng = np.random.default_rng(9)
        buf = ts.VectorReplayBuffer(E * T, E, rew_dtype=torch.float32)
        rew, term, trunc = fill_vec(buf, E, T, rng, 1 / 200, 1 / 1000, rew3=False)
        batch, indices = buf.sample(0)
        v_s = torch.randn(E * T, device=DEV)
        v_n = torch.randn(E * T, device=DEV)
        us = gpu_us(lambda: ts.BasePolicy.compute_episodic_return(batch, buf, indices, v_n, v_s,
                                                                  0.99, 0.95), iters=100)
        # same call with indices that are NOT the canonical tensor: general flat path
        idx2 = indices.clone()
        us_flat = gpu_us(lambda: ts.BasePolicy.compute_episodic_return(
            batch, buf, idx2, v_n, v_s, 0.99, 0.95), iters=100)
        r2 = rew.reshape(E, T).astype(np.float32)
        vs_np, vn_np = v_s.cpu().numpy().reshape(E, T), v_n.cpu().numpy().reshape(E, T)
        t2, c2 = term.reshape(E, T), trunc.reshape(E, T)
        cus = cpu_us(lambda: orc.episodic_return_rows(r2, vs_np, vn_np, t2, c2, 0.99, 0.95),
                     seconds=1.0)
        # ragged sub-buffers (every other env stream is 8 steps short): sample(0) is a listing
        # of runs, not [E, T] rows -> ts_mark_unfinished + the flat kernel without indices
        buf._rings[::2, 1] = T - 8
        buf._rings[::2, 0] = T - 8
        buf._lengths[::2] = T - 8
        buf.last_index[::2] -= 8
        for b in buf.buffers[::2]:
            b._index, b._size = T - 8, T - 8
        buf._version += 1
        rb, ri = buf.sample(0)
        nr = len(rb)
        rvs, rvn = v_s[:nr].contiguous(), v_n[:nr].contiguous()
        assert buf.listing_runs(ri) is not None
        us_runs = gpu_us(lambda: ts.BasePolicy.compute_episodic_return(
            rb, buf, ri, rvn, rvs, 0.99, 0.95), iters=100)
        ri2 = ri.clone()
        us_runs_idx = gpu_us(lambda: ts.BasePolicy.compute_episodic_return(
            rb, buf, ri2, rvn, rvs, 0.99, 0.95), iters=100)
        emit("a1 compute_episodic_return (API)", E=E, T=T, canonical_us=us, general_indices_us=us_flat,
             ragged_listing_us=us_runs, ragged_indexed_us=us_runs_idx,
             cpu_port_us=cus, speedup=cus / us,
             reference_probe_us=2380.0 if (E, T) == (64, 2048) else None)


def row_gae_flat():
    """a1, general form (ts_gae_flat): one flattened sequence, with and without indices.
    Successive launches rotate over 3 distinct input / output sets (no L2 reuse across launches)."""
    ROT = 3
    for E, T in ((4096, 2048), (64, 2048)):
        n = E * T
        sets = []
        for r in range(ROT):
            rng = np.random.default_rng(1 + r)
            sets.append(dict(
                rew=torch.from_numpy(rng.random(n, dtype=np.float32)).to(DEV),
                vs=torch.from_numpy(rng.standard_normal(n, dtype=np.float32)).to(DEV),
                vn=torch.from_numpy(rng.standard_normal(n, dtype=np.float32)).to(DEV),
                term=torch.from_numpy((rng.random(n) < 1 / 200).view(np.uint8)).to(DEV),
                trunc=torch.zeros(n, dtype=torch.uint8, device=DEV),
                idx=torch.arange(n, device=DEV),
                adv=torch.empty(n, device=DEV), ret=torch.empty(n, device=DEV)))
        unf = torch.full((E,), -1, dtype=torch.int64, device=DEV)
        off = torch.arange(E + 1, device=DEV) * T
        lib = L.lib()
        nbytes = lib.ts_gae_flat_workspace_bytes(n)
        ws = torch.zeros(nbytes, dtype=torch.uint8, device=DEV)
        turn = [0]

        def go(with_idx):
            d = sets[turn[0] % ROT]
            turn[0] += 1
            L.check(lib.ts_gae_flat(L.ptr(d["rew"]), L.ptr(d["vs"]), L.ptr(d["vn"]),
                                    L.ptr(d["term"]), L.ptr(d["trunc"]),
                                    L.ptr(d["idx"]) if with_idx else None,
                                    L.ptr(d["term"]) if with_idx else None,
                                    L.ptr(unf) if with_idx else None, L.ptr(off), E, T, n, 0.99,
                                    0.95, 0, 0, 0, L.ptr(d["adv"]), L.ptr(d["ret"]), L.ptr(ws),
                                    nbytes, L.stream()), "flat")
        for with_idx, bpt in ((False, 21), (True, 30)):
            us = gpu_us(lambda: go(with_idx), iters=99)
            emit("a1 gae_flat", E=E, T=T, indices=with_idx, us=us, alg_bytes_per_transition=bpt,
                 GBps=n * bpt / us / 1e3, frac_of_measured_peak=n * bpt / us / 1e3 / PEAK)


def row_gae_flat_sweep():
    """Pipelined-ranges kernel: CTAs per SM x ring stages, two sequence lengths, no indices."""
    lib = L.lib()
    for n in (4096 * 2048, 4000 * 2048, 4000 * 2048 + 7):
        rng = np.random.default_rng(1)
        rew = torch.from_numpy(rng.random(n, dtype=np.float32)).to(DEV)
        vs = torch.from_numpy(rng.standard_normal(n, dtype=np.float32)).to(DEV)
        vn = torch.from_numpy(rng.standard_normal(n, dtype=np.float32)).to(DEV)
        term = torch.from_numpy((rng.random(n) < 1 / 200).view(np.uint8)).to(DEV)
        trunc = torch.zeros(n, dtype=torch.uint8, device=DEV)
        adv, ret = torch.empty(n, device=DEV), torch.empty(n, device=DEV)
        nbytes = lib.ts_gae_flat_workspace_bytes(n)
        ws = torch.zeros(nbytes, dtype=torch.uint8, device=DEV)

        def go(stream=None):
            L.check(lib.ts_gae_flat(L.ptr(rew), L.ptr(vs), L.ptr(vn), L.ptr(term), L.ptr(trunc),
                                    None, None, None, None, 0, 0, n, 0.99, 0.95, 0, 0, 0,
                                    L.ptr(adv), L.ptr(ret), L.ptr(ws), nbytes,
                                    stream if stream is not None else L.stream()), "flat")
        for ctas in (1, 2, 3):
            for stages in (2, 3, 4, 6):
                lib.ts_tune_set(b"gae_flat_ctas", ctas)
                lib.ts_tune_set(b"gae_flat_stages", stages)
                us = gpu_us(go, iters=100)
                gus = graph_us(go)
                emit("a1 gae_flat sweep", N=n, ctas_per_sm=ctas, stages=stages, us=us,
                     graph_us=gus, frac_of_measured_peak=n * 21 / gus / 1e3 / PEAK)
    lib.ts_tune_set(b"gae_flat_ctas", 0)
    lib.ts_tune_set(b"gae_flat_stages", 0)


def row_nstep_index():
    """a2 n-step target and a3/a4 prev/next on the C4 buffer (2^20 transitions, 16 sub-buffers)."""
    E, T = 16, 1 << 16
    rng = np.random.default_rng(3)
    buf = ts.VectorReplayBuffer(E * T, E)
    rew, term, trunc = fill_vec(buf, E, T, rng)
    st = orc.RingState(E * T, E)
    st.rew, st.done, st.terminated = rew, term | trunc, term
    st.last_index, st.lengths = buf.last_index.copy(), buf._lengths.copy()
    for bsz in (32, 4096):
        idx_np = rng.integers(0, E * T, bsz)
        idx = torch.from_numpy(idx_np).to(DEV)
        for D in (1, 51):
            tq = torch.from_numpy(rng.standard_normal((bsz, D)).astype(np.float32)).to(DEV)
            tq_np = tq.cpu().numpy()
            batch = ts.Batch()
            us = gpu_us(lambda: ts.BasePolicy.compute_nstep_return(
                batch, buf, idx, lambda b, t: tq, 0.99, 3))
            cus = cpu_us(lambda: orc.compute_nstep_return(st, idx_np, lambda t: tq_np, 0.99, 3))
            # the two kernels alone (raw C-ABI in a CUDA graph): what the API numbers above hide
            bst = buf._state()
            term_u8, done_u8 = L.as_u8(buf.terminated), L.as_u8(buf.done)
            rew_d = buf.rew.contiguous()
            terminal = torch.empty(bsz, dtype=torch.int64, device=DEV)
            rs = torch.empty(bsz, dtype=torch.float64, device=DEV)
            cf = torch.empty(bsz, dtype=torch.float64, device=DEV)
            outq = torch.empty_like(tq)
            lib = L.lib()

            def kernels(s):
                L.check(lib.ts_nstep_chain(
                    L.ptr(idx), bsz, L.ptr(rew_d), L.dtype_code(rew_d), L.ptr(done_u8),
                    L.ptr(term_u8), L.ptr(bst["offset"]), bst["B"], bst["sub_size"],
                    L.ptr(bst["last_index"]), L.ptr(bst["lengths"]), 0.99, 3, L.ptr(terminal),
                    L.ptr(rs), L.ptr(cf), s), "chain")
                L.check(lib.ts_nstep_combine(L.ptr(tq), L.dtype_code(tq), bsz, D, L.ptr(rs),
                                             L.ptr(cf), L.ptr(outq), s), "combine")
            kus = graph_us(kernels)
            emit("a2 nstep", bsz=bsz, D=D, n_step=3, gpu_us=us, kernels_graph_us=kus,
                 cpu_port_us=cus, speedup=cus / us, kernel_speedup=cus / kus)
        us_n = gpu_us(lambda: buf.next(idx))
        us_p = gpu_us(lambda: buf.prev(idx))
        cus = cpu_us(lambda: st.next(idx_np))
        bst = buf._state()
        done_u8 = L.as_u8(buf.done)
        nout = torch.empty(bsz, dtype=torch.int64, device=DEV)
        lib = L.lib()
        kus = graph_us(lambda s: L.check(lib.ts_next_index(
            L.ptr(idx), bsz, L.ptr(bst["offset"]), bst["B"], bst["sub_size"], L.ptr(done_u8),
            L.ptr(bst["last_index"]), L.ptr(bst["lengths"]), L.ptr(nout), s), "next"))
        emit("a3/a4 next", M=bsz, B=E, gpu_us=us_n, prev_gpu_us=us_p, kernel_graph_us=kus,
             cpu_port_us=cus, speedup=cus / us_n, kernel_speedup=cus / kus)
    # many sub-buffers: the reference's njit loop is O(B * M)
    E2, T2 = 4096, 256
    buf2 = ts.VectorReplayBuffer(E2 * T2, E2)
    rew, term, trunc = fill_vec(buf2, E2, T2, rng)
    st2 = orc.RingState(E2 * T2, E2)
    st2.rew, st2.done, st2.terminated = rew, term | trunc, term
    st2.last_index, st2.lengths = buf2.last_index.copy(), buf2._lengths.copy()
    idx_np = rng.integers(0, E2 * T2, 4096)
    idx = torch.from_numpy(idx_np).to(DEV)
    us = gpu_us(lambda: buf2.next(idx))
    cus = cpu_us(lambda: st2.next(idx_np))
    emit("a3/a4 next", M=4096, B=E2, gpu_us=us, cpu_port_us=cus, speedup=cus / us)


def row_gather():
    """a5: buffer[indices] for Atari-shaped rows (4 x 84 x 84 u8 = 28 224 B)."""
    n_rows, row = 100_000, 4 * 84 * 84
    src = torch.randint(0, 255, (n_rows, row), dtype=torch.uint8, device=DEV)
    rng = np.random.default_rng(4)
    for bsz in (32, 4096):
        idx = torch.from_numpy(rng.integers(0, n_rows, bsz)).to(DEV)
        out = torch.empty(bsz, row, dtype=torch.uint8, device=DEV)
        lib = L.lib()
        us = graph_us(lambda s: L.check(lib.ts_gather_rows(L.ptr(src), n_rows, row, L.ptr(idx), bsz,
                                                           L.ptr(out), s), "gather"))
        emit("a5 gather_rows (graph)", bsz=bsz, row_bytes=row, us=us,
             GBps=2 * bsz * row / us / 1e3, frac_of_measured_peak=2 * bsz * row / us / 1e3 / PEAK)
        assert torch.equal(out, src[idx])


def row_gather_stack():
    """N2: obs + obs_next frame stacks of an Atari-style buffer (stack 4, 84 x 84 u8 frames,
    save_only_last_obs + ignore_obs_next): ts_gather_stack (one launch, k + 1 frames read for 2k
    written) vs the composition of prev / next / row gathers it replaces, and the CPU oracle."""
    B, cap, k, row = 16, 8192, 4, 84 * 84
    total = B * cap
    rng = np.random.default_rng(6)
    frames_np = rng.integers(0, 256, (total, row)).astype(np.uint8)
    offset = np.arange(B + 1, dtype=np.int64) * cap
    lengths = np.full(B, cap, dtype=np.int64)
    last = (offset[:-1] + rng.integers(0, cap, B)).astype(np.int64)
    done = rng.random(total) < 1 / 1000
    d = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(DEV)
    frames, g_off, g_len, g_last, g_done = d(frames_np), d(offset), d(lengths), d(last), \
        d(done.view(np.uint8))
    lib = L.lib()
    for bsz in (32, 256, 4096):
        idx_np = rng.integers(0, total, bsz)
        idx = d(idx_np)
        out = torch.empty(bsz, k, row, dtype=torch.uint8, device=DEV)
        out_next = torch.empty_like(out)

        def fused(s):
            L.check(lib.ts_gather_stack(L.ptr(frames), row, L.ptr(idx), bsz, k, L.ptr(g_off), B,
                                        cap, L.ptr(g_done), L.ptr(g_last), L.ptr(g_len),
                                        L.ptr(out), L.ptr(out_next), s), "gather_stack")

        tmp = torch.empty(bsz, dtype=torch.int64, device=DEV)
        tmp2 = torch.empty(bsz, dtype=torch.int64, device=DEV)
        chain = torch.empty(2, k, bsz, dtype=torch.int64, device=DEV)
        o2 = torch.empty(2, k, bsz, row, dtype=torch.uint8, device=DEV)

        def composed(s):  # what get() did before: k x (row gather + prev), twice, then a stack
            for half in range(2):
                cur = idx
                if half == 1:
                    L.check(lib.ts_next_index(L.ptr(idx), bsz, L.ptr(g_off), B, cap, L.ptr(g_done),
                                              L.ptr(g_last), L.ptr(g_len), L.ptr(tmp2), s), "next")
                    cur = tmp2
                for i in range(k - 1, -1, -1):
                    L.check(lib.ts_gather_rows(L.ptr(frames), total, row, L.ptr(cur), bsz,
                                               L.ptr(o2[half, i]), s), "rows")
                    nxt = chain[half, i]
                    L.check(lib.ts_prev_index(L.ptr(cur), bsz, L.ptr(g_off), B, cap, L.ptr(g_done),
                                              L.ptr(g_last), L.ptr(g_len), L.ptr(nxt), s), "prev")
                    cur = nxt

        us = graph_us(fused)
        cus = graph_us(composed, calls=2)
        ref = cpu_us(lambda: (orc.stack_gather(frames_np, idx_np, k, offset, done, last, lengths),
                              orc.stack_gather(frames_np, idx_np, k, offset, done, last, lengths,
                                               from_next=True)), seconds=1.0)
        fused(L.stream())
        torch.cuda.synchronize()
        want = orc.stack_gather(frames_np, idx_np, k, offset, done, last, lengths, from_next=True)
        assert np.array_equal(out_next.cpu().numpy(), want)
        alg = bsz * 4 * k * row  # 2 x 28 224 read + 2 x 28 224 written per sample
        moved = bsz * (3 * k + 1) * row
        emit("N2 gather_stack obs+obs_next (graph)", bsz=bsz, k=k, row_bytes=row, us=us,
             composed_launches_us=cus, cpu_port_us=ref, alg_GBps=alg / us / 1e3,
             frac_of_measured_peak=alg / us / 1e3 / PEAK, moved_GBps=moved / us / 1e3,
             moved_frac=moved / us / 1e3 / PEAK, samples_per_s=bsz / us * 1e6)


def row_add():
    """N1: one vectorised env step added to a VectorReplayBuffer: add_device (one launch, no host
    read of rew / done) vs the host-bookkeeping add() of this package vs the oracle's ring
    bookkeeping alone on one core (the reference's add is a Python loop over sub-buffers)."""
    for E, obs_dim in ((4096, 27), (64, 27)):
        T = 256
        rng = np.random.default_rng(8)
        step = dict(obs=rng.standard_normal((E, obs_dim)).astype(np.float32),
                    act=rng.standard_normal((E, 8)).astype(np.float32),
                    rew=rng.standard_normal(E), terminated=rng.random(E) < 0.01,
                    truncated=np.zeros(E, dtype=bool),
                    obs_next=rng.standard_normal((E, obs_dim)).astype(np.float32))
        dstep = ts.Batch(**{k: torch.from_numpy(v).to(DEV) for k, v in step.items()})
        hbuf, dbuf = ts.VectorReplayBuffer(E * T, E), ts.VectorReplayBuffer(E * T, E)
        for _ in range(3):
            hbuf.add(ts.Batch(info={}, **step))
            dbuf.add_device(dstep)

        def wall(fn, n=100):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(n):
                fn()
            torch.cuda.synchronize()
            return (time.perf_counter() - t0) / n * 1e6

        host_us = wall(lambda: hbuf.add(ts.Batch(info={}, **step)))
        dev_wall_us = wall(lambda: dbuf.add_device(dstep))
        dev_gpu_us = gpu_us(lambda: dbuf.add_device(dstep), iters=100)
        st = orc.RingState(E * T, E)
        cus = cpu_us(lambda: st.add(step["rew"], step["terminated"], step["truncated"]),
                     seconds=1.0)
        row_bytes = 2 * obs_dim * 4 + 32 + 8 + 3
        emit("N1 add one step", envs=E, bytes_per_transition=row_bytes,
             add_device_wall_us=dev_wall_us, add_device_gpu_us=dev_gpu_us,
             host_add_wall_us=host_us, cpu_port_bookkeeping_us=cus,
             us_per_transition_device=dev_wall_us / E, us_per_transition_host_add=host_us / E,
             reference_probe_us_per_transition=5.5)


def row_ppo():
    """a11: fused PPO loss forward + backward per minibatch (C3: N = 131 072)."""
    N = 131072
    rng = np.random.default_rng(5)
    f = lambda a: torch.from_numpy(a.astype(np.float32)).to(DEV)
    adv, ret, vs = f(rng.standard_normal(N)), f(rng.standard_normal(N)), f(rng.standard_normal(N))
    lo = f(rng.normal(-8, 0.3, N))
    for m in (64, 1024, N):
        mb = torch.from_numpy(rng.permutation(N)[:m]).to(DEV)
        lp = (lo[mb] + 0.1 * torch.randn(m, device=DEV)).requires_grad_(True)
        val = (vs[mb] + 0.3 * torch.randn(m, device=DEV)).requires_grad_(True)

        def fused():
            cl, vl, _ = ts.ppo_loss(lp, val, lo, adv, ret, vs, mb, 0.2, None, True, True)
            (cl + 0.5 * vl).backward()

        def eager():  # the reference's expressions (ppo.py:109-144) in torch eager on the GPU
            a = adv[mb]
            a = (a - a.mean()) / (a.std() + 1e-8)
            ratio = (lp - lo[mb]).exp().float()
            s1, s2 = ratio * a, ratio.clamp(0.8, 1.2) * a
            cl = -torch.min(s1, s2).mean()
            vclip = vs[mb] + (val - vs[mb]).clamp(-0.2, 0.2)
            vl = torch.max((ret[mb] - val).pow(2), (ret[mb] - vclip).pow(2)).mean()
            (cl + 0.5 * vl).backward()

        lpd, vald = lp.detach(), val.detach()
        fwd_us = gpu_us(lambda: ts.ppo_loss(lpd, vald, lo, adv, ret, vs, mb, 0.2, None, True, True),
                        iters=100)  # the call alone: no autograd graph, no backward
        us, eus = gpu_us(fused, iters=100), gpu_us(eager, iters=100)
        lpn, lon = lp.detach().cpu().numpy(), lo[mb].cpu().numpy()
        an, rn = adv[mb].cpu().numpy(), ret[mb].cpu().numpy()
        vn_, valn = vs[mb].cpu().numpy(), val.detach().cpu().numpy()
        cus = cpu_us(lambda: orc.ppo_loss(lpn, lon, an, rn, vn_, valn, 0.2, None, True, True), 1.0)
        emit("a11 ppo_loss fwd+bwd", m=m, fused_gpu_us=us, fused_forward_only_us=fwd_us,
             shim=L.shim() is not None, torch_eager_gpu_us=eus,
             cpu_port_fwd_us=cus, speedup_vs_eager=eus / us)


def row_learn():
    """C3 (BASELINE.json configs[2]): PPOPolicy.learn at Ant shape (obs 27, act 8, 64 x 64 tanh
    MLPs, Independent Normal), wall time per minibatch: this package eager and CUDA-graphed,
    against the reference's learn() arithmetic (ppo.py:99-162: Batch.split fancy-index copies,
    eager elementwise loss, four .item() syncs) in torch eager on the same GPU and on the CPU."""
    D, A = 27, 8

    class Actor(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.mu = torch.nn.Sequential(torch.nn.Linear(D, 64), torch.nn.Tanh(),
                                          torch.nn.Linear(64, 64), torch.nn.Tanh(),
                                          torch.nn.Linear(64, A))
            self.sigma = torch.nn.Parameter(torch.zeros(A))

        def forward(self, obs, state=None, info=None):
            mu = self.mu(obs)
            return (mu, self.sigma.exp().expand_as(mu)), state

    def dist_fn(mu, sigma):
        return torch.distributions.Independent(torch.distributions.Normal(mu, sigma), 1)

    def nets(dev):
        torch.manual_seed(0)
        actor = Actor().to(dev)
        critic = torch.nn.Sequential(torch.nn.Linear(D, 64), torch.nn.Tanh(),
                                     torch.nn.Linear(64, 64), torch.nn.Tanh(),
                                     torch.nn.Linear(64, 1)).to(dev)
        optim = torch.optim.Adam(list(actor.parameters()) + list(critic.parameters()), lr=3e-4)
        return actor, critic, optim

    def data(n, dev):
        g = torch.Generator().manual_seed(1)
        r = lambda *sh: torch.randn(*sh, generator=g)
        v_s = r(n)
        return dict(obs=r(n, D).to(dev), act=r(n, A).to(dev), v_s=v_s.to(dev),
                    adv=r(n).to(dev), returns=(v_s + 0.3 * r(n)).to(dev),
                    logp_old=(-11.0 + 0.3 * r(n)).to(dev))

    def reference_learn(d, actor, critic, optim, bs, dev):
        n = d["obs"].shape[0]
        params = list(actor.parameters()) + list(critic.parameters())
        perm = np.random.permutation(n)
        out = []
        for lo in range(0, n, bs):
            idx = perm[lo:lo + bs]
            mb = {k: v[idx] for k, v in d.items()}  # Batch.split: a fancy-index copy per key
            dist = dist_fn(*actor(mb["obs"])[0])
            a = mb["adv"]
            a = (a - a.mean()) / (a.std() + 1e-8)
            ratio = (dist.log_prob(mb["act"]) - mb["logp_old"]).exp().float()
            ratio = ratio.reshape(ratio.size(0), -1).transpose(0, 1)
            s1, s2 = ratio * a, ratio.clamp(0.8, 1.2) * a
            clip_loss = -torch.min(s1, s2).mean()
            value = critic(mb["obs"]).flatten()
            v_clip = mb["v_s"] + (value - mb["v_s"]).clamp(-0.2, 0.2)
            vf_loss = torch.max((mb["returns"] - value).pow(2),
                                (mb["returns"] - v_clip).pow(2)).mean()
            ent = dist.entropy().mean()
            loss = clip_loss + 0.25 * vf_loss - 0.0 * ent
            optim.zero_grad()
            loss.backward()
            torch.nn.utils.clip_grad_norm_(params, max_norm=0.5)
            optim.step()
            out.append((clip_loss.item(), vf_loss.item(), ent.item(), loss.item()))
        return out

    def wall_us(fn, reps=3):
        fn()
        best = 1e30
        for _ in range(reps):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            fn()
            torch.cuda.synchronize()
            best = min(best, time.perf_counter() - t0)
        return best * 1e6

    for bs, count in ((64, 256), (1024, 128)):
        n = bs * count
        res = {}
        for mode in ("eager", "cuda_graph"):
            actor, critic, optim = nets(DEV)
            pol = ts.PPOPolicy(actor, critic, optim, dist_fn, eps_clip=0.2, value_clip=True,
                               vf_coef=0.25, ent_coef=0.0, max_grad_norm=0.5,
                               cuda_graph=(mode == "cuda_graph"))
            batch = ts.Batch(**data(n, DEV))
            res[mode] = wall_us(lambda: pol.learn(batch, bs, 1)) / count
        actor, critic, optim = nets(DEV)
        d = data(n, DEV)
        res["ref_gpu"] = wall_us(lambda: reference_learn(d, actor, critic, optim, bs, DEV)) / count
        cpu = torch.device("cpu")
        actor, critic, optim = nets(cpu)
        nc = bs * 32
        dc = data(nc, cpu)
        res["ref_cpu"] = wall_us(lambda: reference_learn(dc, actor, critic, optim, bs, cpu),
                                 reps=2) / 32
        emit("C3 PPOPolicy.learn per minibatch", minibatch=bs, minibatches=count,
             this_eager_us=res["eager"], this_cuda_graph_us=res["cuda_graph"],
             reference_arithmetic_torch_gpu_us=res["ref_gpu"],
             reference_arithmetic_torch_cpu_us=res["ref_cpu"], cpu_threads=torch.get_num_threads(),
             speedup_graph_vs_ref_gpu=res["ref_gpu"] / res["cuda_graph"],
             speedup_graph_vs_ref_cpu=res["ref_cpu"] / res["cuda_graph"])


def row_dqn():
    """C4 (BASELINE.json configs[3]) through the public API: DQNPolicy.update(bsz, buffer) on a
    PrioritizedVectorReplayBuffer of 2^20 Atari-shaped transitions (84 x 84 u8 frames, stack 4,
    obs_next from next(), 16 sub-buffers), n-step 3, Nature CNN.  Wall time per segment of the
    update pipeline (policy/base.py:288-315) and of the networks alone."""
    E, T, A = 16, 1 << 16, 6
    buf = ts.PrioritizedVectorReplayBuffer(E * T, E, alpha=0.6, beta=0.4, stack_num=4,
                                           ignore_obs_next=True, save_only_last_obs=True)
    rng = np.random.default_rng(4)
    for t in range(2):
        buf.add(ts.Batch(obs=np.zeros((E, 4, 84, 84), np.uint8), act=np.zeros(E, np.int64),
                         rew=np.zeros(E), terminated=np.zeros(E, bool),
                         truncated=np.zeros(E, bool), info={}))
    n = E * T
    term = rng.random(n) < 1 / 1000
    trunc = (rng.random(n) < 1 / 10000) & ~term
    buf._meta.rew.copy_(torch.from_numpy(rng.choice(np.array([-1.0, 0.0, 1.0]), n,
                                                    p=[0.05, 0.9, 0.05])))
    buf._meta.act.copy_(torch.from_numpy(rng.integers(0, A, n)))
    buf._meta.terminated.copy_(torch.from_numpy(term))
    buf._meta.truncated.copy_(torch.from_numpy(trunc))
    buf._meta.done.copy_(torch.from_numpy(term | trunc))
    buf._meta.obs.random_(0, 256)
    buf._rings[:, 0], buf._rings[:, 1] = 0, T
    buf._lengths[:] = T
    buf.last_index[:] = np.arange(E) * T + T - 1
    buf._version += 1
    for b in buf.buffers:
        b._index, b._size = 0, T
    buf.update_weight(torch.arange(n, device=DEV), torch.randn(n, device=DEV))

    class Net(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.c = torch.nn.Sequential(
                torch.nn.Conv2d(4, 32, 8, 4), torch.nn.ReLU(), torch.nn.Conv2d(32, 64, 4, 2),
                torch.nn.ReLU(), torch.nn.Conv2d(64, 64, 3, 1), torch.nn.ReLU(),
                torch.nn.Flatten(), torch.nn.Linear(3136, 512), torch.nn.ReLU(),
                torch.nn.Linear(512, A))

        def forward(self, obs, state=None, info=None):
            return self.c(obs.float() * (1.0 / 255)), state

    def wall(fn, iters):
        for _ in range(5):
            fn()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(iters):
            fn()
        torch.cuda.synchronize()
        return (time.perf_counter() - t0) / iters * 1e6

    for bsz in (32, 256):
        torch.manual_seed(0)
        net = Net().to(DEV)
        optim = torch.optim.Adam(net.parameters(), lr=1e-4)
        pol = ts.DQNPolicy(net, optim, 0.99, estimation_step=3, target_update_freq=500)
        np.random.seed(0)
        iters = 100
        total = wall(lambda: pol.update(bsz, buf), iters)
        torch.manual_seed(0)
        net_g = Net().to(DEV)
        pol_g = ts.DQNPolicy(net_g, torch.optim.Adam(net_g.parameters(), lr=1e-4), 0.99,
                             estimation_step=3, target_update_freq=500, cuda_graph=True)
        graphed = wall(lambda: pol_g.update(bsz, buf), iters)
        seg = {}
        state = {}

        def s_sample():
            state["b"], state["i"] = buf.sample(bsz)

        def s_process():
            state["b"].weight = state["w"].clone()
            pol.process_fn(state["b"], buf, state["i"])

        def s_learn():
            state["b"].weight = state["w"].clone()
            pol.learn(state["b"])

        def s_post():
            pol.post_process_fn(state["b"], buf, state["i"])

        seg["sample"] = wall(s_sample, iters)
        state["w"] = state["b"].weight
        seg["process_fn"] = wall(s_process, iters)
        seg["learn"] = wall(s_learn, iters)
        seg["post_process_fn"] = wall(s_post, iters)
        obs = state["b"].obs
        tgt = torch.randn(bsz, device=DEV)
        act = state["b"].act.long().reshape(-1, 1)

        def nets_only():  # target forward x2 (double DQN) + online forward/backward + Adam
            with torch.no_grad():
                net(obs)[0]
                pol.model_old(obs)[0]
            optim.zero_grad()
            q = net(obs)[0].gather(1, act).squeeze(1)
            ((tgt - q) ** 2).mean().backward()
            optim.step()

        seg["networks_only"] = wall(nets_only, iters)
        emit("C4 DQNPolicy.update through the API", bsz=bsz, update_wall_us=total,
             update_cuda_graph_wall_us=graphed, speedup_graph=total / graphed,
             **{k + "_us": v for k, v in seg.items()},
             data_path_share=1.0 - seg["networks_only"] / total)


if __name__ == "__main__":
    torch.cuda.set_device(0)
    for kv in [x for x in os.environ.get("TS_TUNE", "").split(",") if x]:  # A/B of tuning knobs
        key, val = kv.split("=")
        L.check(L.lib().ts_tune_set(key.encode(), int(val)), "tune " + key)
    which = sys.argv[1:] or ["gae_api", "gae_flat", "nstep_index", "gather", "gather_stack", "add", "ppo", "learn"]
    for w in which:
        globals()["row_" + w]()
