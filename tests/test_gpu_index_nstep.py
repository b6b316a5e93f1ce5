"""prev / next / unfinished_index / sample_indices(0) and n-step returns on the GPU vs the
golden fixtures (reference outputs) and the CPU oracle.  Integer outputs: bit-exact."""
import numpy as np
import pytest
import torch

from conftest import load_golden
from helpers import INDEX_SPECS, NSTEP_SPECS, fill
from oracle import oracle as orc

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("spec", INDEX_SPECS, ids=[s[0] for s in INDEX_SPECS])
def test_vector_buffer_index_goldens(spec):
    import tianshou_fork_b200 as ts
    name, n_env, total, steps, ragged = spec
    c = load_golden("index")[name]
    buf = ts.VectorReplayBuffer(total, n_env)
    fill(buf, ts.Batch, 200 + INDEX_SPECS.index(spec), steps, n_env, 0.15, 0.05, ragged)
    assert np.array_equal(buf._extend_offset, c["offset"])
    assert np.array_equal(buf.done.cpu().numpy(), c["done"])
    assert np.array_equal(buf.last_index, c["last_index"])
    assert np.array_equal(buf._lengths, c["lengths"])
    assert np.array_equal([b._index for b in buf.buffers], c["ring_index"])
    q = c["query"]
    # numpy in -> numpy out
    assert np.array_equal(buf.prev(q), c["prev"]) and np.array_equal(buf.next(q), c["next"])
    # tensor in -> tensor out
    qd = torch.from_numpy(q).cuda()
    assert np.array_equal(buf.prev(qd).cpu().numpy(), c["prev"])
    assert np.array_equal(buf.next(qd).cpu().numpy(), c["next"])
    # scalar in -> scalar out, list accepted (test/base/test_buffer.py:893-895)
    assert buf.prev(int(q[5])) == c["prev"][5] and np.ndim(buf.next(-1)) == 0
    assert np.array_equal(buf.next(list(q[:7])), c["next"][:7])
    assert np.array_equal(buf.unfinished_index(), c["unfinished"])
    assert np.array_equal(buf.sample_indices(0).cpu().numpy(), c["all_indices"])
    assert len(buf) == c["lengths"].sum()
    # the children see the same rings (views of the manager's storage)
    k = n_env - 1
    child = buf.buffers[k]
    assert len(child) == c["lengths"][k]
    local = np.arange(len(child))
    if len(child):
        glob = buf.next(local + c["offset"][k]) - c["offset"][k]
        assert np.array_equal(child.next(local), glob)


def test_unequal_manager_goldens():
    import tianshou_fork_b200 as ts
    c = load_golden("index")["m_unequal"]
    buf = ts.ReplayBufferManager([ts.ReplayBuffer(5), ts.ReplayBuffer(9), ts.ReplayBuffer(3),
                                  ts.ReplayBuffer(16)])
    fill(buf, ts.Batch, 250, 12, 4, 0.2, 0.1, True)
    assert buf._equal_sub == 0  # binary-search path
    assert np.array_equal(buf.done.cpu().numpy(), c["done"])
    assert np.array_equal(buf.prev(c["query"]), c["prev"])
    assert np.array_equal(buf.next(c["query"]), c["next"])
    assert np.array_equal(buf.unfinished_index(), c["unfinished"])
    assert np.array_equal(buf.sample_indices(0).cpu().numpy(), c["all_indices"])
    with pytest.raises(NotImplementedError):
        buf.update(ts.ReplayBuffer(4))


@pytest.mark.parametrize("name,size,steps", [("s_partial", 16, 9), ("s_wrap", 16, 41)])
def test_single_ring_goldens(name, size, steps):
    import tianshou_fork_b200 as ts
    c = load_golden("index")[name]
    rng = np.random.default_rng(260 + steps)
    buf = ts.ReplayBuffer(size)
    for _ in range(steps):
        term = rng.random() < 0.2
        buf.add(ts.Batch(obs=np.zeros(2), act=0, rew=rng.standard_normal(), terminated=term,
                         truncated=(not term) and rng.random() < 0.1, obs_next=np.zeros(2),
                         info={}))
    assert len(buf) == int(c["size"]) and buf._index == int(c["ring_index"])
    assert np.array_equal(buf.done.cpu().numpy()[:len(buf)], c["done"][:len(buf)])
    assert np.array_equal(buf.prev(c["query"]), c["prev"])
    assert np.array_equal(buf.next(c["query"]), c["next"])
    assert np.array_equal(buf.unfinished_index(), c["unfinished"])
    assert np.array_equal(buf.sample_indices(0).cpu().numpy(), c["all_indices"])
    with pytest.raises(IndexError):
        buf[size * 2]  # noqa  (test/base/test_buffer.py:127-128)
    with pytest.raises(AssertionError):
        buf.rew = 1  # reserved key (base.py:89-93)
    with pytest.raises(AttributeError):
        buf.no_such_key  # noqa


def test_neighbours_random_differential_large():
    """4096 random queries on a 64 x 2048 buffer vs the oracle, equal and unequal layouts."""
    from tianshou_fork_b200 import _lib as L
    rng = np.random.default_rng(9)
    for equal in (True, False):
        B = 64
        sizes = np.full(B, 2048) if equal else rng.integers(1, 4096, B)
        offset = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
        cap = int(offset[-1])
        lengths = np.minimum(sizes, rng.integers(0, 5000, B)).astype(np.int64)
        lengths[3] = 0
        last = (offset[:-1] + rng.integers(0, np.maximum(lengths, 1))).astype(np.int64)
        done = rng.random(cap) < 0.01
        q = rng.integers(-cap, 2 * cap, 4096).astype(np.int64)
        d = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
        g_q, g_off, g_last, g_len = d(q), d(offset), d(last), d(lengths)
        g_done = d(done.view(np.uint8))
        out = torch.empty_like(g_q)
        sub = int(sizes[0]) if equal else 0
        for fn, ref in ((L.lib().ts_prev_index, orc.prev_index),
                        (L.lib().ts_next_index, orc.next_index)):
            L.check(fn(L.ptr(g_q), q.size, L.ptr(g_off), B, sub, L.ptr(g_done), L.ptr(g_last),
                       L.ptr(g_len), L.ptr(out), L.stream()), "neighbour")
            assert np.array_equal(out.cpu().numpy(), ref(q, offset, done, last, lengths))


@pytest.mark.parametrize("spec", NSTEP_SPECS, ids=[s[0] for s in NSTEP_SPECS])
def test_nstep_goldens(spec):
    """compute_nstep_return vs the reference's output (make_golden.gen_nstep)."""
    import tianshou_fork_b200 as ts
    name, n_env, total, steps, n_step, D = spec
    c = load_golden("nstep")[name]
    buf = ts.VectorReplayBuffer(total, n_env)
    fill(buf, ts.Batch, 300 + NSTEP_SPECS.index(spec), steps, n_env, 0.1, 0.05,
         name == "n5_partial")
    assert np.array_equal(buf.done.cpu().numpy(), c["done"])
    assert np.array_equal(buf.rew.cpu().numpy(), c["rew"])
    table = torch.from_numpy(c["table"]).cuda()
    seen = {}

    def target_q_fn(buffer, indices):
        seen["terminal"] = indices
        t = table[indices]
        return t if D > 1 else t.flatten()

    indice = torch.from_numpy(c["indice"]).cuda()
    out = ts.BasePolicy.compute_nstep_return(ts.Batch(), buf, indice, target_q_fn, gamma=0.97,
                                             n_step=n_step)
    assert out.returns.shape == c["returns"].shape and out.returns.dtype == torch.float32
    # the f64 Horner sum is evaluated in the reference's order -> identical after the f32 cast
    assert np.array_equal(out.returns.cpu().numpy(), c["returns"]), name
    # terminal == indices[-1] of the reference's next() chain (integer: bit-exact)
    st = orc.RingState(total, n_env)
    st.done, st.last_index, st.lengths = c["done"], c["last_index"], c["lengths"]
    term = c["indice"]
    for _ in range(n_step - 1):
        term = st.next(term)
    assert np.array_equal(seen["terminal"].cpu().numpy(), term)
    # numpy indices -> target_q_fn gets numpy, result still on the target's device
    def target_q_np(buffer, indices):
        assert isinstance(indices, np.ndarray)
        return torch.from_numpy(c["table"][indices] if D > 1 else c["table"][indices].ravel())
    b2 = ts.Batch(weight=np.ones(len(c["indice"])))
    out2 = ts.BasePolicy.compute_nstep_return(b2, buf, c["indice"], target_q_np, gamma=0.97,
                                              n_step=n_step)
    assert not out2.returns.is_cuda and np.array_equal(out2.returns.numpy(), c["returns"])
    assert isinstance(out2.weight, torch.Tensor) and out2.weight.dtype == torch.float32
    with pytest.raises(AssertionError):
        ts.BasePolicy.compute_nstep_return(ts.Batch(), buf, indice, target_q_fn, rew_norm=True)


def test_nstep_known_answers_from_reference_tests():
    """test/base/test_returns.py:170-230: ring of 10, wrap-around, n = 1, 2, 10."""
    import tianshou_fork_b200 as ts
    buf = ts.ReplayBuffer(10)
    for i in range(12):
        buf.add(ts.Batch(obs=0, act=0, rew=i + 1, terminated=i % 4 == 3, truncated=False,
                         obs_next=0, info={}))
    indices = np.array([2, 3, 4, 5, 6, 7, 8, 9, 0, 1])

    def target_q_fn(buffer, idx):  # test_returns.py:137-140: minus the NEXT reward
        nxt = buffer.next(idx)
        return -buffer.rew[torch.as_tensor(nxt, device="cuda")].float()

    for n, expect in [(1, [2.6, 4, 4.4, 5.3, 6.2, 8, 8, 8.9, 9.8, 12]),
                      (2, [3.4, 4, 5.53, 6.62, 7.8, 8, 9.89, 10.98, 12.2, 12]),
                      (10, [3.4, 4, 5.678, 6.78, 7.8, 8, 10.122, 11.22, 12.2, 12])]:
        r = ts.BasePolicy.compute_nstep_return(ts.Batch(), buf, indices, target_q_fn, gamma=.1,
                                               n_step=n).returns
        assert np.allclose(r.cpu().numpy().ravel(), expect), n


def test_seeded_sample_indices_vs_reference():
    """sample_indices(batch_size > 0) draws from the global np.random stream in the reference's
    call order (base.py:276-305, manager.py:163-192): a seeded run returns the reference's
    indices (tests/golden/sample.npz, make_golden.gen_sample)."""
    import tianshou_fork_b200 as ts
    g = load_golden("sample")
    specs = [("vec_partial", 4, 40, 7, {}), ("vec_wrapped", 3, 30, 25, {}),
             ("vec_stack_avail", 4, 64, 11, dict(stack_num=3, sample_avail=True)),
             ("vec_one_empty", 5, 50, 6, {})]
    for name, nb, total, steps, kw in specs:
        buf = ts.VectorReplayBuffer(total, nb, **kw)
        rng = np.random.default_rng(950)
        for t in range(steps):
            ids = np.arange(nb if name != "vec_one_empty" else nb - 1)
            m = ids.size
            term = rng.random(m) < 0.15
            buf.add(ts.Batch(obs=rng.standard_normal((m, 2)).astype(np.float32),
                             act=rng.integers(0, 3, m), rew=rng.standard_normal(m),
                             terminated=term, truncated=np.zeros(m, bool),
                             obs_next=rng.standard_normal((m, 2)).astype(np.float32), info={}),
                    buffer_ids=ids)
        np.random.seed(31)
        for i, bsz in enumerate((1, 7, 64)):
            got = buf.sample_indices(bsz).cpu().numpy()
            assert np.array_equal(got, g[name][f"draw{i}"]), (name, bsz)
    for name, size, steps, kw in [("ring", 16, 9, {}), ("ring_wrapped", 16, 41, {}),
                                  ("ring_stack_avail", 32, 20,
                                   dict(stack_num=4, sample_avail=True))]:
        buf = ts.ReplayBuffer(size, **kw)
        rng = np.random.default_rng(951)
        for t in range(steps):
            buf.add(ts.Batch(obs=rng.standard_normal(2).astype(np.float32),
                             act=int(rng.integers(3)), rew=float(rng.standard_normal()),
                             terminated=bool(rng.random() < 0.2), truncated=False,
                             obs_next=rng.standard_normal(2).astype(np.float32), info={}))
        np.random.seed(32)
        for i, bsz in enumerate((1, 5, 40)):
            got = buf.sample_indices(bsz).cpu().numpy()
            assert np.array_equal(got, g[name][f"draw{i}"]), (name, bsz)


def test_from_data_buffer_matches_added_buffer():
    """ReplayBuffer.from_data (base.py:108-132): a buffer built over existing arrays behaves
    like one filled by add() -- same neighbours, same gathered rows."""
    import tianshou_fork_b200 as ts
    rng = np.random.default_rng(8)
    n = 40
    data = dict(obs=rng.standard_normal((n, 3)).astype(np.float32), act=rng.integers(0, 4, n),
                rew=rng.standard_normal(n), terminated=rng.random(n) < 0.2,
                truncated=np.zeros(n, bool),
                obs_next=rng.standard_normal((n, 3)).astype(np.float32))
    added = ts.ReplayBuffer(n)
    for i in range(n):
        added.add(ts.Batch(info={}, **{k: v[i] for k, v in data.items()}))
    built = ts.ReplayBuffer.from_data(done=data["terminated"] | data["truncated"], **data)
    assert len(built) == n
    q = np.arange(n)
    # from_data leaves last_index at 0 like the reference (add() left it at n - 1): compare
    # away from the slots whose neighbour test looks at either of them
    assert np.array_equal(built.prev(q)[2:], added.prev(q)[2:])
    assert np.array_equal(built.next(q)[1:-1], added.next(q)[1:-1])
    b1, b2 = built[q], added[q]
    for k in ("obs", "act", "rew", "terminated", "done", "obs_next"):
        assert torch.equal(b1[k].to(b2[k].dtype), b2[k]), k
