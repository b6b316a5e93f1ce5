#!/usr/bin/env python
"""Multi-GPU check of the sharded PER path on real GPUs (run under torchrun, one rank per GPU):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29533 scripts/sharded_check.py [p2p|nccl]

Every rank owns 2^16 leaves; all ranks also build the reference's ONE global tree with the CPU
oracle (test infrastructure) and check, over chained sample -> update steps with duplicated
indices, that the union of the ranks' samples is bit-identical to the global tree's, that the
normalised importance weights agree to 1e-12, and time the step."""
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    mode = sys.argv[1] if len(sys.argv) > 1 else "p2p"
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device(f"cuda:{local}")
    dist.init_process_group("nccl", device_id=dev)
    import tianshou_fork_b200 as ts
    from oracle import oracle as orc

    cap, nbuf, bsz = 1 << 16, 4, 4096
    total = world * cap
    buf = ts.ShardedPrioritizedVectorReplayBuffer(total, world * nbuf, alpha=0.6, beta=0.4,
                                                  exchange=mode)
    assert buf.exchange == mode, (buf.exchange, mode)
    rng = np.random.default_rng(5)  # the same stream on every rank
    leaves = (np.abs(rng.standard_normal(total)) + 1e-7) ** 0.6
    glob = orc.PER(total, 0.6, 0.4)
    glob.weight[np.arange(total)] = leaves
    buf.weight[torch.arange(cap, device=dev)] = torch.from_numpy(
        leaves[rank * cap:(rank + 1) * cap]).to(dev)
    buf._lengths[:] = cap // nbuf
    buf.sync_roots()
    for step in range(6):
        u = rng.random(bsz)
        want = glob.sample_indices(u)
        want_w = glob.batch_weight(want)
        idx = buf.sample_with_uniforms(torch.from_numpy(u).to(dev))
        w = buf._cached_weight[1]
        buf.check_errors()
        own = (idx >= 0)
        owned = own.to(torch.int64)
        dist.all_reduce(owned)
        assert bool((owned == 1).all()), "every query owned by exactly one rank"
        gi = buf.to_global(idx).cpu().numpy()
        m = own.cpu().numpy()
        assert np.array_equal(gi[m], want[m]), (rank, step)
        assert np.allclose(w.cpu().numpy()[m], want_w[m], rtol=1e-12), (rank, step)
        td = rng.standard_normal(bsz).astype(np.float32)
        gidx = want.copy()
        gidx[:300] = gidx[300:600]
        glob.update_weight(gidx, td)
        loc = np.where((gidx >= rank * cap) & (gidx < (rank + 1) * cap), gidx - rank * cap, -1)
        buf.update_weight(torch.from_numpy(loc).to(dev), torch.from_numpy(td).to(dev))
        got_l = buf.weight._value[cap:].cpu().numpy()
        ref_l = glob.weight._value[total + rank * cap: total + (rank + 1) * cap]
        assert np.allclose(got_l, ref_l, rtol=6e-7, atol=0)
        ch = np.nonzero(got_l != ref_l)[0]  # re-anchor on the oracle's f32 powers
        if ch.size:
            buf.weight[torch.from_numpy(ch).to(dev)] = torch.from_numpy(ref_l[ch]).to(dev)
        buf.sync_roots()
        buf.check_errors()
    # ---- a batch of several update passes (4096 * world draws): the sampler's owner list ---------
    big = 4096 * world + 123
    for step in range(3):
        u = rng.random(big)
        want = glob.sample_indices(u)
        idx = buf.sample_with_uniforms(torch.from_numpy(u).to(dev))
        own = idx >= 0
        gi = buf.to_global(idx).cpu().numpy()
        m = own.cpu().numpy()
        assert np.array_equal(gi[m], want[m]), (rank, "big", step)
        if mode == "p2p":
            assert buf._owned_src is not None and int(buf._owned[2]) == int(own.sum()), "owner list"
        td = rng.standard_normal(big).astype(np.float32)
        glob.update_weight(want, td)
        buf.update_weight(idx, torch.from_numpy(td).to(dev))  # the sampler's own tensor: listed scan
        got_l = buf.weight._value[cap:].cpu().numpy()
        ref_l = glob.weight._value[total + rank * cap: total + (rank + 1) * cap]
        assert np.allclose(got_l, ref_l, rtol=6e-7, atol=0), (rank, "big", step)
        ch = np.nonzero(got_l != ref_l)[0]
        if ch.size:
            buf.weight[torch.from_numpy(ch).to(dev)] = torch.from_numpy(ref_l[ch]).to(dev)
        buf.sync_roots()
        buf.check_errors()
    # prev / next / n-step on the sharded buffer in this exchange mode (ReplayBuffer._state must
    # stay callable next to the exchange's own state words)
    buf._meta = ts.Batch(rew=torch.zeros(cap, dtype=torch.float64, device=dev),
                         done=torch.zeros(cap, dtype=torch.bool, device=dev),
                         terminated=torch.zeros(cap, dtype=torch.bool, device=dev))
    q = torch.arange(0, cap, 37, device=dev)
    nxt = buf.next(q).cpu().numpy()
    qn, sub = q.cpu().numpy(), cap // nbuf
    want_n = np.where((qn + 1) % sub == 0, qn - sub + 1, qn + 1)
    want_n = np.where(np.isin(qn, np.asarray(buf.last_index)), qn, want_n)  # the newest slot stays
    assert np.array_equal(nxt, want_n), "next"
    r = ts.BasePolicy.compute_nstep_return(ts.Batch(), buf, q,
                                           lambda b, i: torch.ones(i.numel(), device=dev), 0.5, 2)
    stays = torch.from_numpy(np.isin(qn, np.asarray(buf.last_index))).to(dev)
    want_r = torch.where(stays, 0.5, 0.25).to(r.returns.dtype)  # rew 0: target * gamma^steps
    assert torch.equal(r.returns.flatten(), want_r), "n-step on a sharded buffer"
    # ---- timing: eager steps (host launches included), identical draws on every rank ----------
    R = 8
    us = torch.from_numpy(rng.random((R, bsz))).to(dev)
    tds = torch.from_numpy(rng.standard_normal((R, bsz)).astype(np.float32)).to(dev)

    def step(r):
        i = buf.sample_with_uniforms(us[r])
        buf.update_weight(i, tds[r])

    for r in range(R):
        step(r)
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 200
    t0 = time.perf_counter()
    e0.record()
    for i in range(n):
        step(i % R)
    e1.record()
    torch.cuda.synchronize()
    wall = (time.perf_counter() - t0) / n * 1e6
    buf.check_errors()
    t = torch.tensor([e0.elapsed_time(e1) * 1e3 / n], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        print(f"sharded PER check ok: world={world} exchange={mode} "
              f"step={float(t):.1f} us (device, max over ranks) wall={wall:.1f} us", flush=True)
    dist.barrier()
    sys.stdout.flush()
    os._exit(0)  # destroy_process_group() has hung after symmetric-memory use (profiles/r02_per.md)


if __name__ == "__main__":
    main()
