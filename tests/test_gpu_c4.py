"""BASELINE.json configs[3] at FULL size on the GPU: a PrioritizedVectorReplayBuffer of 2^20 slots
in 16 wrapped rings, batches 32 / 256 / 4096, n-step 3 with D = 1 and 51 -- against outputs of
the reference itself (tests/golden/c4.npz, make_golden.gen_c4) and the CPU oracle."""
import hashlib

import numpy as np
import pytest
import torch

import helpers as H
from conftest import load_golden
from oracle import oracle as orc

pytestmark = pytest.mark.gpu


def _sha(a):
    return np.frombuffer(hashlib.sha256(np.ascontiguousarray(a).tobytes()).digest(), np.uint8)


@pytest.fixture(scope="module")
def c4():
    """The device buffer in the reference's final state.  The rollout (65 536 + 1234 steps x 16
    envs) goes in through ``add`` for the first and the last 300 steps -- allocation, wrap-around
    and bookkeeping are the class's own -- and as one block copy in between (storage order is
    a pure function of the step number)."""
    import tianshou_fork_b200 as ts
    rew, term, trunc = H.c4_stream()
    E, sub = H.C4_E, H.C4_SUB
    steps = rew.shape[0]
    buf = ts.PrioritizedVectorReplayBuffer(E * sub, E, alpha=0.6, beta=0.4)
    zeros = np.zeros(E, np.uint8)

    def add(t):
        buf.add(ts.Batch(obs=zeros, act=zeros, rew=rew[t], terminated=term[t],
                         truncated=trunc[t], obs_next=zeros, info={}))

    head, tail = 300, 300
    for t in range(head):
        add(t)
    # steps [head, steps - tail): rings advance by (steps - tail - head) without per-step calls
    mid = steps - tail - head
    buf._rings[:, 0] = (buf._rings[:, 0] + mid) % sub
    buf._rings[:, 1] = np.minimum(buf._rings[:, 1] + mid, sub)
    buf._lengths[:] = buf._rings[:, 1]
    buf._version += 1
    for t in range(steps - tail, steps):
        add(t)
    s_rew, s_term, s_trunc, ring_index, last = H.c4_final_storage(rew, term, trunc)
    d = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
    buf._meta.rew.copy_(d(s_rew))
    buf._meta.terminated.copy_(d(s_term))
    buf._meta.truncated.copy_(d(s_trunc))
    buf._meta.done.copy_(d(s_term | s_trunc))
    assert np.all(buf._rings[:, 0] == ring_index) and np.array_equal(buf.last_index, last)
    buf.weight[torch.arange(E * sub, device="cuda")] = d(H.c4_leaves())
    st = orc.RingState(E * sub, E)
    st.rew, st.terminated, st.truncated, st.done = s_rew, s_term, s_trunc, s_term | s_trunc
    st.index[:], st.size[:], st.lengths[:] = ring_index, sub, sub
    st.last_index = last
    return buf, st, load_golden("c4")["c4"]


def test_c4_state_and_neighbours(c4):
    buf, st, g = c4
    assert np.array_equal(_sha(buf.rew.cpu().numpy()), g["rew_sha"])
    assert np.array_equal(_sha(buf.done.cpu().numpy()), g["done_sha"])
    assert np.array_equal(buf.last_index, g["last_index"])
    assert np.array_equal(buf._lengths, g["lengths"])
    assert np.array_equal(buf.unfinished_index(), g["unfinished"])
    assert buf.weight.reduce() == float(g["root0"])  # 2^20 leaves, every level bit-exact
    assert np.array_equal(buf.prev(g["q"]), g["prev"])
    assert np.array_equal(buf.next(g["q"]), g["next"])
    q = torch.from_numpy(g["q"]).cuda()
    assert np.array_equal(buf.prev(q).cpu().numpy(), g["prev"])
    assert np.array_equal(buf.next(q).cpu().numpy(), g["next"])


def test_c4_sample_nstep_update_chain(c4):
    """sample -> buffer[idx].weight -> compute_nstep_return (D = 1, 51) -> update_weight for batch
    32, 256, 4096, chained on one tree like make_golden.gen_c4 ran it on the reference."""
    import tianshou_fork_b200 as ts
    buf, st, g = c4
    size = H.C4_E * H.C4_SUB
    rng = np.random.default_rng(4002)
    rng.integers(0, size, 4096)  # the generator drew q first
    for bsz in H.C4_BATCHES:
        np.random.seed(2000 + bsz)
        idx = buf.sample_indices(bsz)  # np.random.rand(bsz) on the host like prio.py:61
        assert np.array_equal(idx.cpu().numpy(), g[f"b{bsz}/idx"]), bsz  # bit-exact
        batch = buf[idx]
        assert np.allclose(batch.weight.cpu().numpy(), g[f"b{bsz}/weight"], rtol=1e-12)
        assert np.array_equal(batch.rew.cpu().numpy(), st.rew[g[f"b{bsz}/idx"]])
        for D in H.C4_DIMS:
            def target_q_fn(buffer, terminal, D=D):
                t = torch.from_numpy(H.c4_target_q(terminal.cpu().numpy(), D)).cuda()
                return t if D > 1 else t.flatten()

            r = ts.BasePolicy.compute_nstep_return(ts.Batch(), buf, idx, target_q_fn, gamma=0.99,
                                                   n_step=3).returns.cpu().numpy()
            assert r.dtype == np.float32 and r.shape == (bsz, D)
            assert np.array_equal(r[:64], g[f"b{bsz}/d{D}/returns_head"])
            assert np.array_equal(_sha(r), g[f"b{bsz}/d{D}/returns_sha"]), (bsz, D)
            want = orc.compute_nstep_return(st, g[f"b{bsz}/idx"],
                                            lambda t: H.c4_target_q(t, D), 0.99, 3)
            assert np.array_equal(r, want.astype(np.float32).reshape(bsz, D))
        td = rng.standard_normal(bsz).astype(np.float32)
        idx_u = g[f"b{bsz}/idx"].copy()
        k = bsz // 20
        idx_u[:k] = idx_u[k:2 * k]  # 5 % duplicated indices: the last occurrence wins
        buf.update_weight(torch.from_numpy(idx_u).cuda(), torch.from_numpy(td).cuda())
        uniq = np.unique(idx_u)
        # f32 `power` of NumPy (prio.py:87) is up to ~2.6 f32 ulp off the exact value; ours is
        # the f64 power rounded once: float compare, then re-anchor on the
        # reference's leaves so the root and the NEXT batch's indices test bit-exactness again
        got = buf.weight[uniq]
        assert np.allclose(got, g[f"b{bsz}/leaves"], rtol=6e-7, atol=0)
        assert np.array_equal([buf._max_prio, buf._min_prio], g[f"b{bsz}/prio"])
        buf.weight[torch.from_numpy(uniq).cuda()] = torch.from_numpy(g[f"b{bsz}/leaves"]).cuda()
        assert buf.weight.reduce() == float(g[f"b{bsz}/root"])
    buf.check_errors()


@pytest.mark.parametrize("bsz", [32, 256, 4096])
def test_c4_nstep_abi_random_batches_vs_oracle(c4, bsz):
    """ts_nstep_chain + ts_nstep_combine on uniformly random indices (every slot class: ring seam,
    unfinished tails, terminated / truncated steps), D = 51, against the oracle."""
    import tianshou_fork_b200 as ts
    buf, st, _ = c4
    rng = np.random.default_rng(bsz)
    idx = rng.integers(0, H.C4_E * H.C4_SUB, bsz)
    idx[: H.C4_E] = st.last_index                       # the unfinished tails themselves
    idx[H.C4_E: 2 * H.C4_E] = st.prev(st.last_index)    # one step before them
    for n in (1, 3, 5):
        def target_q_fn(buffer, terminal):
            return torch.from_numpy(H.c4_target_q(terminal.cpu().numpy(), 51)).cuda()

        r = ts.BasePolicy.compute_nstep_return(ts.Batch(), buf, torch.from_numpy(idx).cuda(),
                                               target_q_fn, gamma=0.99, n_step=n).returns
        want = orc.compute_nstep_return(st, idx, lambda t: H.c4_target_q(t, 51), 0.99, n)
        assert np.array_equal(r.cpu().numpy(), want.astype(np.float32)), n
