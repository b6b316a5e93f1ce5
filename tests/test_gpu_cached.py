"""CachedReplayBuffer on device storage: the known answers of the reference's own test
(test/base/test_buffer.py:904-1031) and seeded add streams replayed against goldens written by
the reference (tests/golden/cached.npz, make_golden.gen_cached).  All outputs are integer or
byte copies: bit-exact."""
import pickle

import numpy as np
import pytest
import torch

from conftest import load_golden

pytestmark = pytest.mark.gpu


def _np(x):
    return x.cpu().numpy() if isinstance(x, torch.Tensor) else np.asarray(x)


def _step(ts, vals, term, ids=None):
    n = len(vals)
    return ts.Batch(obs=np.array(vals, float), act=np.array(vals, float),
                    rew=np.array(vals, float), terminated=np.array(term, bool),
                    truncated=np.zeros(n, bool))


def test_known_answers_of_the_reference_test():
    import tianshou_fork_b200 as ts
    buf = ts.CachedReplayBuffer(ts.ReplayBuffer(10), 4, 5)
    assert _np(buf.sample_indices(0)).tolist() == []
    ptr, ep_rew, ep_len, ep_idx = buf.add(_step(ts, [1], [0]), buffer_ids=[1])
    idx = buf.sample_indices(0)
    assert _np(idx).tolist() == [15]
    assert _np(buf.prev(idx)).tolist() == [15] and _np(buf.next(idx)).tolist() == [15]
    assert (ptr, ep_idx, ep_len, ep_rew) == ([15], [15], [0], [0.0])
    ptr, ep_rew, ep_len, ep_idx = buf.add(_step(ts, [2], [1]), buffer_ids=[3])
    idx = buf.sample_indices(0)
    assert _np(idx).tolist() == [0, 15]
    assert _np(buf.prev(idx)).tolist() == [0, 15] and _np(buf.next(idx)).tolist() == [0, 15]
    assert (ptr, ep_idx, ep_len, ep_rew) == ([0], [0], [1], [2.0])
    assert _np(buf.unfinished_index()).tolist() == [15]
    ptr, ep_rew, ep_len, ep_idx = buf.add(_step(ts, [3, 4], [0, 1]), buffer_ids=[3, 1])
    assert ep_len.tolist() == [0, 2] and ep_rew.tolist() == [0.0, 5.0]
    assert ptr.tolist() == [25, 2] and ep_idx.tolist() == [25, 1]
    obs = np.zeros(buf.maxsize)
    obs[[0, 1, 2, 15, 16, 25]] = [2, 1, 4, 1, 4, 3]
    assert np.array_equal(_np(buf.obs), obs)
    assert _np(buf.unfinished_index()).tolist() == [25]
    idx = buf.sample_indices(0)
    assert _np(idx).tolist() == [0, 1, 2, 25]
    assert _np(buf.done)[_np(idx)].tolist() == [1, 0, 1, 0]
    assert _np(buf.prev(idx)).tolist() == [0, 1, 1, 25]
    assert _np(buf.next(idx)).tolist() == [0, 2, 2, 25]
    np.random.seed(0)
    counts = np.bincount(_np(buf.sample_indices(10000)), minlength=26)
    assert counts[[0, 1, 2, 25]].min() > 2000 and counts.sum() == 10000


def test_main_buffer_of_size_zero_and_vector_rewards():
    import tianshou_fork_b200 as ts
    buf = ts.CachedReplayBuffer(ts.ReplayBuffer(0, sample_avail=True), 4, 5)
    data, rew = np.zeros(4), np.ones([4, 4])
    for term in ([0, 0, 1, 1], [0, 0, 0, 0], [1, 1, 1, 1], [0, 0, 0, 0]):
        buf.add(ts.Batch(obs=data, act=data, rew=rew, terminated=np.array(term, bool),
                         truncated=np.zeros(4, bool)))
    ptr, ep_rew, ep_len, ep_idx = buf.add(
        ts.Batch(obs=data, act=data, rew=rew, terminated=np.array([0, 1, 0, 1], bool),
                 truncated=np.zeros(4, bool)))
    assert ptr.tolist() == [1, -1, 11, -1] and ep_idx.tolist() == [0, -1, 10, -1]
    assert ep_len.tolist() == [0, 2, 0, 2]
    assert np.array_equal(ep_rew, np.array([data, data + 2, data, data + 2]))
    want_done = np.zeros(20, bool)
    want_done[[2, 6, 7, 16]] = True
    assert np.array_equal(_np(buf.done), want_done)
    idx = buf.sample_indices(0)
    assert _np(idx).tolist() == [0, 1, 10, 11]
    assert _np(buf.prev(idx)).tolist() == [0, 0, 10, 10]
    assert _np(buf.next(idx)).tolist() == [1, 1, 11, 11]


@pytest.mark.parametrize("name", ["roomy", "main_wraps", "long_episodes"])
def test_seeded_stream_vs_reference(name):
    import tianshou_fork_b200 as ts
    from helpers import CACHED_SPECS
    g = load_golden("cached")[name]
    _, main, n_cached, ep_len, steps, p_done, ragged = next(s for s in CACHED_SPECS
                                                           if s[0] == name)
    buf = ts.CachedReplayBuffer(ts.ReplayBuffer(main), n_cached, ep_len)
    rng = np.random.default_rng(900)
    rets = []
    for t in range(steps):
        ids = np.arange(n_cached)
        if ragged:
            ids = ids[rng.random(n_cached) < 0.7]
            if ids.size == 0:
                continue
        m = ids.size
        term = rng.random(m) < p_done
        trunc = (rng.random(m) < p_done / 3) & ~term
        b = ts.Batch(obs=rng.standard_normal((m, 2)).astype(np.float32),
                     act=rng.integers(0, 4, m), rew=rng.standard_normal(m), terminated=term,
                     truncated=trunc, obs_next=rng.standard_normal((m, 2)).astype(np.float32),
                     info={})
        ptr, ep_rew, ep_l, ep_idx = buf.add(b, buffer_ids=ids)
        rets.append(np.stack([ptr, ep_rew, ep_l, ep_idx], 1).astype(np.float64))
        if t == steps // 2:  # a pickle round trip in the middle of the stream changes nothing
            buf = pickle.loads(pickle.dumps(buf))
    assert np.array_equal(np.concatenate(rets), g["add_returns"])
    assert np.array_equal(buf._lengths, g["lengths"])
    assert np.array_equal(buf.last_index, g["last_index"]) and len(buf) == int(g["len"])
    allidx = np.arange(buf.maxsize)
    assert np.array_equal(_np(buf.sample_indices(0)), g["all"])
    assert np.array_equal(_np(buf.prev(allidx)), g["prev"])
    assert np.array_equal(_np(buf.next(allidx)), g["next"])
    assert np.array_equal(_np(buf.unfinished_index()), g["unfinished"])
    for k in ("obs", "act", "rew", "done", "terminated", "obs_next"):
        assert np.array_equal(_np(buf._meta[k]), g[k]), k
