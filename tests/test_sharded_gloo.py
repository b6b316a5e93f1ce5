"""world_size-2 / -4 `gloo` tests (CPU) of the sharded-PER host protocol: the
shard layout, the 24-byte root exchange of ``tianshou_fork_b200.data.sharded`` and the claim
it rests on -- G power-of-two-aligned shard trees + the top levels rebuilt from their roots
reproduce the reference's single global tree bit for bit.  The trees themselves are the CPU
oracle here (the CUDA kernels are checked against the same oracle in test_gpu_sharded.py)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import oracle as orc


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank: int, world: int, port: int, total: int, buffers: int, q):
    try:
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        dist.init_process_group("gloo", rank=rank, world_size=world)
        from tianshou_fork_b200.data.sharded import exchange_roots, is_exact_layout, shard_layout
        size, per_rank, cap, offset = shard_layout(total, buffers, world, rank)
        assert is_exact_layout(cap, world) and offset == rank * cap
        assert size * buffers == total and per_rank * world == buffers
        # the global leaf array is the same on every rank (same seed); each keeps its slice
        rng = np.random.default_rng(7)
        leaves = (np.abs(rng.standard_normal(total)) + 1e-7) ** 0.6
        glob = orc.SegTree(total)
        glob[np.arange(total)] = leaves
        mine = orc.SegTree(cap)
        mine[np.arange(cap)] = leaves[offset:offset + cap]
        for step in range(3):
            mine3 = torch.tensor([mine.reduce(), 1.0 + rank + step, 1.0 / (1 + rank + step)],
                                 dtype=torch.float64)
            roots3 = exchange_roots(mine3, world).numpy()
            assert roots3.shape == (world, 3) and roots3[rank, 0] == mine.reduce()
            assert roots3[:, 1].max() == world + step and roots3[:, 2].min() == 1.0 / (world + step)
            u = rng.random(2048)  # identical draws on every rank
            total_p, shard, v = orc.sharded_route(u, roots3[:, 0])
            assert total_p == glob.reduce()  # pairwise top == the global root, bit for bit
            want = glob.get_prefix_sum_idx(u * glob.reduce())
            own = shard == rank
            got = mine.get_prefix_sum_idx(v[own].copy()) + offset
            assert np.array_equal(got, want[own]), (rank, step)
            # every query is owned by exactly one rank
            owned = torch.from_numpy(own.astype(np.int64))
            dist.all_reduce(owned)
            assert bool((owned == 1).all())
            # an update on every rank, mirrored on the global tree: local positions only
            k = 300
            pos = rng.integers(0, total, k)
            val = rng.random(k)
            glob[pos] = val
            sel = (pos >= offset) & (pos < offset + cap)
            if sel.any():
                mine[pos[sel] - offset] = val[sel]
            assert np.array_equal(mine._value[cap:], glob._value[total + offset:total + offset + cap])
        dist.destroy_process_group()
        q.put((rank, "ok"))
    except Exception as e:  # surface the failure in the parent
        import traceback
        q.put((rank, traceback.format_exc()))


@pytest.mark.parametrize("world,total,buffers", [(2, 1 << 12, 8), (4, 1 << 13, 16)])
def test_sharded_protocol_gloo(world, total, buffers):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, total, buffers, q))
             for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(60)
    assert all(msg == "ok" for _, msg in results), results


def test_shard_layout_rejects_unaligned():
    from tianshou_fork_b200.data.sharded import is_exact_layout, shard_layout
    size, per_rank, cap, off = shard_layout(8_000_000, 128, 8, 3)
    assert (size, per_rank, cap, off) == (62500, 16, 1_000_000, 3_000_000)
    assert not is_exact_layout(cap, 8)  # 10^6 per shard straddles subtrees
    size, per_rank, cap, off = shard_layout(8 << 20, 128, 8, 7)
    assert cap == 1 << 20 and off == 7 << 20 and is_exact_layout(cap, 8)


def _dp_worker(rank: int, world: int, port: int, q):
    try:
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        dist.init_process_group("gloo", rank=rank, world_size=world)
        from tianshou_fork_b200.policy.base import BasePolicy
        torch.manual_seed(0)  # identical replicas
        net = torch.nn.Sequential(torch.nn.Linear(5, 7), torch.nn.Tanh(), torch.nn.Linear(7, 2))
        x_all = torch.randn(world * 6, 5)
        y_all = torch.randn(world * 6, 2)
        pol = BasePolicy()
        pol.set_data_parallel(True)
        # rank-local loss on this rank's rows, then the averaged gradient
        (net(x_all[rank * 6:(rank + 1) * 6]) - y_all[rank * 6:(rank + 1) * 6]).pow(2).mean().backward()
        pol._sync_grads(net.parameters())
        got = [p.grad.clone() for p in net.parameters()]
        net.zero_grad()
        (net(x_all) - y_all).pow(2).mean().backward()  # one learner over every rank's rows
        for g, p in zip(got, net.parameters()):
            assert torch.allclose(g, p.grad, rtol=1e-5, atol=1e-7)
        pol.set_data_parallel(False)
        before = [p.grad.clone() for p in net.parameters()]
        pol._sync_grads(net.parameters())  # disabled: untouched, no collective
        assert all(torch.equal(a, p.grad) for a, p in zip(before, net.parameters()))
        dist.destroy_process_group()
        q.put((rank, "ok"))
    except Exception:
        import traceback
        q.put((rank, traceback.format_exc()))


def test_data_parallel_gradient_average_gloo():
    """SURVEY 8e, learner row: per-rank gradients summed over the group and divided by its
    size equal the gradient of one learner over all ranks' (equal-sized) minibatches."""
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_dp_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(60)
    assert all(msg == "ok" for _, msg in results), results
