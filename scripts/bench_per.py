"""PER leg of bench.py (BASELINE.json configs[3] on one GPU, configs[4] on N GPUs).

One 2^20-leaf f64 sum tree per GPU.  A step = sample (descent + importance weights + batch-max
normalisation) + update_weight on the sampled indices, 8 steps per CUDA graph, 50 replays.

N == 1: ts_per_sample_lazy + ts_per_update_lazy (the update stops at its range roots, the
        sampling cluster rebuilds the top of the tree in shared memory); the eager pair
        ts_per_sample + ts_per_update is timed beside it.
N  > 1: the reference's ONE global tree of N * 2^20 leaves, sharded, the shard
        records and the weight maximum exchanged INSIDE the kernels over NVLink peer memory.
        Two legs: "strong" (global batch 4096, the reference's C5 config) and "weak" (global
        batch 4096 * N: every rank routes all draws and descends ~4096 of them).  Before the
        timing, rank 0 checks the first steps against the oracle's single global tree.
"""
from __future__ import annotations

import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def note(msg):
    print(f"[bench_per r{os.environ.get('RANK', '0')} {time.strftime('%H:%M:%S')}] {msg}",
          file=sys.stderr, flush=True)

PER_LEAVES, PER_BATCH, PER_ALPHA, PER_BETA = 1 << 20, 4096, 0.6, 0.4
PER_ALG_BYTES_PER_SAMPLE = 664
R = 8  # distinct (uniforms, td) sets per graph


def shard_leaves(rank: int) -> np.ndarray:
    rng = np.random.default_rng(100 + rank)
    return (np.abs(rng.standard_normal(PER_LEAVES)) + 1e-7) ** PER_ALPHA


class Shard:
    """Raw device state of one 2^20-leaf shard (what the buffer classes hold)."""

    def __init__(self, torch, L, lib, dev, rank):
        size = bound = PER_LEAVES
        self.tree = torch.zeros(2 * bound, dtype=torch.float64, device=dev)
        self.stamp = torch.zeros(size, dtype=torch.int32, device=dev)
        self.counter = torch.zeros(4, dtype=torch.int32, device=dev)
        self.err = torch.zeros(1, dtype=torch.int32, device=dev)
        self.prio = torch.ones(2, dtype=torch.float64, device=dev)
        leaves = torch.from_numpy(shard_leaves(rank)).to(dev)
        all_idx = torch.arange(size, device=dev)
        st = torch.cuda.current_stream().cuda_stream
        L.check(lib.ts_segtree_update(self.tree.data_ptr(), size, bound, all_idx.data_ptr(),
                                      leaves.data_ptr(), L.TS_F64, 1, size, self.stamp.data_ptr(),
                                      self.counter.data_ptr(), self.err.data_ptr(), st),
                "ts_segtree_update")


def time_graph(torch, dev, step, barrier, replays=50):
    """8 steps captured into one CUDA graph; returns (ms of `replays` replays, steps)."""
    side = torch.cuda.Stream(dev)
    with torch.cuda.stream(side):
        for r in range(R):
            step(r, side.cuda_stream)
    side.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph, stream=side):
        for r in range(R):
            step(r, side.cuda_stream)
    for _ in range(5):
        graph.replay()
    barrier()
    stream = torch.cuda.current_stream()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(replays):
        graph.replay()
    e1.record(stream)
    barrier()
    return e0.elapsed_time(e1), replays * R


def run_single(torch, L, lib, dev, barrier, max_over_ranks):
    sh = Shard(torch, L, lib, dev, 0)
    size = bound = PER_LEAVES
    shared = np.random.default_rng(2)
    u = torch.from_numpy(shared.random((R, PER_BATCH))).to(dev)
    td = torch.from_numpy(shared.standard_normal((R, PER_BATCH)).astype(np.float32)).to(dev)
    idx = torch.empty(R, PER_BATCH, dtype=torch.int64, device=dev)
    w = torch.empty(R, PER_BATCH, dtype=torch.float64, device=dev)
    t = sh.tree.data_ptr()

    def eager_pair(r, s):
        L.check(lib.ts_per_sample(t, bound, u[r].data_ptr(), PER_BATCH, sh.prio.data_ptr(),
                                  PER_BETA, 1, idx[r].data_ptr(), w[r].data_ptr(), None,
                                  sh.err.data_ptr(), s), "ts_per_sample")
        L.check(lib.ts_per_update(t, size, bound, idx[r].data_ptr(), td[r].data_ptr(), L.TS_F32,
                                  PER_BATCH, PER_ALPHA, sh.prio.data_ptr(), sh.stamp.data_ptr(),
                                  sh.counter.data_ptr(), sh.err.data_ptr(), s), "ts_per_update")

    def lazy_pair(r, s):
        L.check(lib.ts_per_sample_lazy(t, bound, u[r].data_ptr(), PER_BATCH, sh.prio.data_ptr(),
                                       PER_BETA, 1, idx[r].data_ptr(), w[r].data_ptr(),
                                       sh.err.data_ptr(), s), "ts_per_sample_lazy")
        L.check(lib.ts_per_update_lazy(t, size, bound, idx[r].data_ptr(), td[r].data_ptr(),
                                       L.TS_F32, PER_BATCH, PER_ALPHA, sh.prio.data_ptr(),
                                       sh.stamp.data_ptr(), sh.counter.data_ptr(),
                                       sh.err.data_ptr(), s), "ts_per_update_lazy")

    ms_e, steps_e = time_graph(torch, dev, eager_pair, barrier)
    ms, steps = time_graph(torch, dev, lazy_pair, barrier)
    L.check(lib.ts_segtree_finish_top(t, bound, torch.cuda.current_stream().cuda_stream), "finish")
    assert int(sh.err.item()) == 0, "PER range flag raised"
    us = ms * 1e3 / steps
    return {
        "value": PER_BATCH * steps / (ms * 1e-3), "us_per_step": us,
        "eager_pair_us_per_step": ms_e * 1e3 / steps_e, "gpu_launches_per_step": 2,
        "scaling": "n/a (1 GPU)",
        "how": "8 steps per CUDA graph; ts_per_sample_lazy + ts_per_update_lazy (the update ends "
               "at its range roots, the sampling cluster rebuilds the 7 levels above them in "
               "shared memory and writes them back); eager_pair = ts_per_sample + ts_per_update",
    }


def _parity_check(torch, dist, L, lib, dev, rank, world, sh, step_one, u_all, td_all, idx, steps=3,
                  reread=False):
    """Rank 0 compares the first `steps` draws of the sharded path with the oracle's SINGLE global
    tree of world * 2^20 leaves (manager.py:209-214) on identical uniforms.  After every
    update_weight the oracle's touched leaves are re-anchored on the GPU's (NumPy's f32 `power`
    is not bit-reproducible across libms), then its root must equal the
    sharded root bit for bit, so every later draw tests exactness again."""
    from oracle import oracle as orc  # checker only (task statement 4: bench N > 1 parity)
    M = u_all.shape[1]
    ref = None
    if reread:  # the shards have been updated since they were filled: gather their leaves as they are
        mine_l = sh.tree[PER_LEAVES:].clone()
        all_l = [torch.empty_like(mine_l) for _ in range(world)]
        dist.all_gather(all_l, mine_l)
        start = torch.cat(all_l).cpu().numpy() if rank == 0 else None
    elif rank == 0:
        start = np.concatenate([shard_leaves(r) for r in range(world)])
    if rank == 0:
        ref = orc.SegTree(world * PER_LEAVES)
        ref[np.arange(world * PER_LEAVES)] = start
    st = torch.cuda.current_stream().cuda_stream
    mism = 0
    for k in range(steps):
        step_one(k, st, sample=True, update=False)
        mine = idx[k].clone()
        glob = torch.where(mine >= 0, mine + rank * PER_LEAVES, torch.full_like(mine, -1))
        gathered = [torch.empty_like(glob) for _ in range(world)]
        dist.all_gather(gathered, glob)
        merged = torch.stack(gathered).max(0).values
        owners = (torch.stack(gathered) >= 0).sum(0)
        assert int(owners.min()) == 1 and int(owners.max()) == 1, "one owner per draw"
        step_one(k, st, sample=False, update=True)
        # the leaves this rank's update wrote, for the oracle
        local = torch.where(mine >= 0, mine, torch.zeros_like(mine))
        vals = torch.where(mine >= 0, sh.tree[PER_LEAVES + local], torch.full_like(
            local, -1.0, dtype=torch.float64))
        gv = [torch.empty_like(vals) for _ in range(world)]
        dist.all_gather(gv, vals)
        root = sh.tree[1:2].clone()
        roots = [torch.empty_like(root) for _ in range(world)]
        dist.all_gather(roots, root)
        if rank == 0:
            want = ref.get_prefix_sum_idx(u_all[k].cpu().numpy() * ref.reduce())
            got = merged.cpu().numpy()
            mism += int(np.count_nonzero(want != got))
            new_leaf = torch.stack(gv).max(0).values.cpu().numpy()
            # last occurrence wins (segtree.py:40-49): NumPy assignment order does the same
            ref[got] = new_leaf
            level = [float(r.item()) for r in roots]
            while len(level) > 1:  # the global tree's pairwise order above the shard roots
                level = [level[i] + level[i + 1] for i in range(0, len(level), 2)]
            assert level[0] == ref.reduce(), "sharded root != oracle root after re-anchoring"
    out = torch.tensor([mism], dtype=torch.int64, device=dev)
    dist.broadcast(out, 0)
    return {"parity": "bit-exact" if int(out.item()) == 0 else f"{int(out.item())} mismatches",
            "checked_steps": steps, "checked_draws": steps * M,
            "against": "oracle SegTree over all shards' leaves (one global tree), same uniforms"}


def run_sharded(torch, dist, L, lib, dev, rank, world, barrier, max_over_ranks):
    import torch.distributed._symmetric_memory as symm_mem
    sh = Shard(torch, L, lib, dev, rank)
    size = bound = PER_LEAVES
    slots = symm_mem.empty(16 * world, dtype=torch.float64, device=dev)
    slots.zero_()
    hdl = symm_mem.rendezvous(slots, dist.group.WORLD)
    peers = int(hdl.buffer_ptrs_dev)
    state = torch.zeros(2, dtype=torch.int64, device=dev)
    own_list = torch.zeros(4 + PER_BATCH * world, dtype=torch.int32, device=dev)  # weak leg's owner list
    torch.cuda.synchronize()
    dist.barrier()
    t = sh.tree.data_ptr()
    st0 = torch.cuda.current_stream().cuda_stream
    L.check(lib.ts_per_publish(t, sh.prio.data_ptr(), state.data_ptr(), peers, world, rank, st0),
            "ts_per_publish")
    legs = {}
    parity = weak_parity = None
    for leg, M in (("strong", PER_BATCH), ("weak", PER_BATCH * world)):
        shared = np.random.default_rng(2 + M)  # the same draws on every rank
        u = torch.from_numpy(shared.random((R, M))).to(dev)
        td = torch.from_numpy(shared.standard_normal((R, M)).astype(np.float32)).to(dev)
        idx = torch.empty(R, M, dtype=torch.int64, device=dev)
        w = torch.empty(R, M, dtype=torch.float64, device=dev)

        def step(r, s, sample=True, update=True, M=M, u=u, td=td, idx=idx, w=w):
            # batches of more than one update pass hand the owned positions from the sampler to
            # the update (what ShardedPrioritizedVectorReplayBuffer does for them)
            own = own_list.data_ptr() if M > PER_BATCH and os.environ.get("TS_PER_NO_OWNED") != "1" else None
            if sample:
                L.check(lib.ts_per_sample_p2p_owned(
                    t, bound, u[r].data_ptr(), M, slots.data_ptr(), peers, state.data_ptr(), world,
                    rank, sh.prio.data_ptr(), PER_BETA, 1, idx[r].data_ptr(), w[r].data_ptr(), own,
                    sh.err.data_ptr(), s), "ts_per_sample_p2p_owned")
            if update:
                L.check(lib.ts_per_update_publish_owned(
                    t, size, bound, idx[r].data_ptr(), td[r].data_ptr(), L.TS_F32, M, PER_ALPHA,
                    sh.prio.data_ptr(), sh.stamp.data_ptr(), sh.counter.data_ptr(),
                    sh.err.data_ptr(), state.data_ptr(), peers, world, rank, own, s),
                    "ts_per_update_publish_owned")

        note(f"sharded leg {leg}, global batch {M}")
        # both legs start from the leaves the oracle is built from only in the strong leg; the
        # weak leg's check re-reads this rank's leaves first
        check = _parity_check(torch, dist, L, lib, dev, rank, world, sh, step, u, td, idx,
                              steps=3 if leg == "strong" else 2, reread=leg != "strong")
        if leg == "strong":
            parity = check
        else:
            weak_parity = check["parity"]
        note("parity done, timing")
        ms, steps = time_graph(torch, dev, step, barrier)
        ms = max_over_ranks(ms)
        owned = (idx >= 0).sum().to(torch.float64)
        dist.all_reduce(owned)
        assert int(owned.item()) == R * M, "every query must be owned by exactly one rank"
        legs[leg] = {"global_batch": M, "us_per_step": ms * 1e3 / steps,
                     "value": M * steps / (ms * 1e-3), "unit": "samples/s"}
    assert int(sh.err.item()) == 0, "PER range / peer-timeout flag raised"
    return {
        "value": legs["strong"]["value"], "us_per_step": legs["strong"]["us_per_step"],
        "strong": legs["strong"], "weak": {**legs["weak"], "parity": weak_parity,
                                            "owner_list": os.environ.get("TS_PER_NO_OWNED") != "1"},
        **parity, "gpu_launches_per_step": 2,
        "scaling": "strong (global batch 4096 over all shards) in `value`; weak (4096 per rank) "
                   "in `weak`",
        "how": "8 steps per CUDA graph; ts_per_sample_p2p + ts_per_update_publish: shard records "
               "{root, max_prio, min_prio} and the weight maximum are exchanged INSIDE the "
               "kernels over NVLink peer memory (st.release.sys / ld.acquire.sys on symmetric "
               "memory), no NCCL call on the step",
    }


def per_e2e(torch, dist, dev, rank, world, max_over_ranks):
    """The PER step through the reference-facing classes with HOST arrays: uniforms drawn by
    np.random on the host (prio.py:61) and copied in, sampled indices and importance weights read
    back to the host, host td-errors copied in for update_weight -- every step, inside the timed
    region (wall clock; every step ends in a blocking device-to-host read).  N > 1: the sharded
    class, identical host draws on every rank, indices local to the rank (-1: another rank's)."""
    try:
        import tianshou_fork_b200 as ts
        E = 16
        if world == 1:
            buf = ts.PrioritizedVectorReplayBuffer(PER_LEAVES, E, alpha=PER_ALPHA, beta=PER_BETA)
            api = "PrioritizedVectorReplayBuffer"
        else:
            buf = ts.ShardedPrioritizedVectorReplayBuffer(PER_LEAVES * world, E * world,
                                                          alpha=PER_ALPHA, beta=PER_BETA)
            api = f"ShardedPrioritizedVectorReplayBuffer (exchange: {buf.exchange})"
        for _ in range(2):  # allocate storage; the ring bookkeeping is then declared full
            buf.add(ts.Batch(obs=np.zeros((E, 1), np.float32), act=np.zeros(E, np.int64),
                             rew=np.zeros(E), terminated=np.zeros(E, bool),
                             truncated=np.zeros(E, bool),
                             obs_next=np.zeros((E, 1), np.float32), info={}))
        T = PER_LEAVES // E
        buf._rings[:, 0], buf._rings[:, 1] = 0, T
        buf._lengths[:] = T
        buf.last_index[:] = np.arange(E) * T + T - 1
        buf._version += 1
        rng = np.random.default_rng(3 + rank)
        buf.update_weight(torch.arange(PER_LEAVES, device=dev),
                          torch.from_numpy(rng.standard_normal(PER_LEAVES).astype(np.float32)
                                           ).to(dev))
        tds = [rng.standard_normal(PER_BATCH).astype(np.float32) for _ in range(8)]
        np.random.seed(12345)  # the same host draws on every rank

        def step(i):
            if world == 1:
                # host draws up, indices + IS weights down as one packed copy (prio.py:59-64, 91-102)
                idx_h, w_h = buf.sample_indices_numpy(PER_BATCH)
            else:
                idx = buf.sample_indices(PER_BATCH)        # host uniforms -> device descent
                w = buf._cached_weight[1]                  # IS weights of the same launch
                idx_h, w_h = idx.cpu().numpy(), w.cpu().numpy()  # the learner's host copies
            buf.update_weight(idx_h, tds[i % 8])           # host td-errors -> device tree
            return idx_h, w_h

        for i in range(20):
            step(i)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        steps = 300
        t0 = time.perf_counter()
        for i in range(steps):
            step(i)
        torch.cuda.synchronize()
        dt = max_over_ranks(time.perf_counter() - t0)
        buf.check_errors()
        return {"value": PER_BATCH * steps / dt, "unit": "samples/s",
                "us_per_step": dt / steps * 1e6,
                "h2d_bytes_per_step": PER_BATCH * (8 + 8 + 4),
                "d2h_bytes_per_step": PER_BATCH * (8 + 8),
                "api": api + (".sample_indices_numpy (host draws up, indices + importance weights "
                              "down in one packed copy) + update_weight(NumPy, NumPy) (one packed "
                              "pinned upload)" if world == 1 else
                              ".sample_indices + importance weights + update_weight with NumPy "
                              "arrays in and out")}
    except Exception as e:  # reported, not fatal -- on ONE GPU.  With peers, a rank that drops out
        if world > 1:       # of a collective sequence would leave them waiting: fail the run
            raise
        return {"error": f"{type(e).__name__}: {e}"}


def run_per(torch, dist, L, lib, dev, rank, world, barrier, max_over_ranks, peak):
    if world == 1:
        res = run_single(torch, L, lib, dev, barrier, max_over_ranks)
    else:
        res = run_sharded(torch, dist, L, lib, dev, rank, world, barrier, max_over_ranks)
    how = res.pop("how")
    us = res["us_per_step"]
    achieved = PER_BATCH * PER_ALG_BYTES_PER_SAMPLE / (us * 1e-6) / 1e9
    out = {"metric": "PER samples/s", "unit": "samples/s", **res,
           "workload": f"{world} shard(s) x 2^20-leaf f64 sum tree, alpha 0.6 beta 0.4; step = "
                       "sample (descent + IS weights + batch-max norm) + update_weight (|td|+eps, "
                       "pow, last-writer table in shared memory, per-range rebuild; programmatic "
                       "dependent launch on both kernels); " + how,
           "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                        "frac": achieved / peak, "traffic": None,
                        "alg_bytes_per_step": PER_BATCH * PER_ALG_BYTES_PER_SAMPLE,
                        "note": "latency-bound: ~9 dependent L2 round trips of ~0.5 us per step "
                                "plus two launch boundaries, not bytes; the tree is L2-resident "
                                "(ncu: ~0 B DRAM read per launch, profiles/r01c_per.md)"}}
    note("per e2e through the buffer class")
    out["e2e"] = per_e2e(torch, dist, dev, rank, world, max_over_ranks)
    return out
