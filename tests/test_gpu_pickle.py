"""Pickle round trips of device-resident buffers (reference: base.py:81-87,
test/base/test_buffer.py:597-632): contents, bookkeeping, the sum tree and the behaviour
after loading (prev / next / add / sample) must be those of the original."""
import pickle

import numpy as np
import pytest
import torch

from helpers import golden_stream

pytestmark = pytest.mark.gpu


def _same_storage(a, b):
    for k in ("obs", "act", "rew", "terminated", "truncated", "done", "obs_next"):
        if k in a._meta.__dict__:
            assert torch.equal(a._meta[k], b._meta[k]), k


def test_pickle_single_ring_like_reference_test():
    import tianshou_fork_b200 as ts
    size = 100
    vbuf = ts.ReplayBuffer(size, stack_num=2)
    pbuf = ts.PrioritizedReplayBuffer(size, 0.6, 0.4)
    for i in range(4):
        vbuf.add(ts.Batch(obs=np.array([i]), act=0, rew=1.0, terminated=0, truncated=0))
    for i in range(5):
        pbuf.add(ts.Batch(obs=np.array([i]), act=2, rew=1.0, terminated=0, truncated=0))
    pbuf.update_weight(np.array([1, 3]), np.array([0.5, 2.0]))
    _vbuf = pickle.loads(pickle.dumps(vbuf))
    _pbuf = pickle.loads(pickle.dumps(pbuf))
    assert len(_vbuf) == len(vbuf) and torch.equal(_vbuf.act, vbuf.act)
    assert len(_pbuf) == len(pbuf) and torch.equal(_pbuf.act, pbuf.act)
    assert _vbuf.stack_num == vbuf.stack_num
    q = np.arange(len(pbuf))
    assert np.array_equal(np.asarray(_pbuf.weight[q]), np.asarray(pbuf.weight[q]))
    assert _pbuf._max_prio == pbuf._max_prio and _pbuf._min_prio == pbuf._min_prio
    assert np.array_equal(_vbuf.prev(np.arange(4)), vbuf.prev(np.arange(4)))
    # the loaded buffer keeps working
    _vbuf.add(ts.Batch(obs=np.array([9]), act=1, rew=2.0, terminated=1, truncated=0))
    vbuf.add(ts.Batch(obs=np.array([9]), act=1, rew=2.0, terminated=1, truncated=0))
    _same_storage(_vbuf, vbuf)


@pytest.mark.parametrize("device_adds", [False, True])
def test_pickle_vector_buffer_roundtrip(device_adds):
    import tianshou_fork_b200 as ts
    n_env, total = 4, 48
    buf = ts.PrioritizedVectorReplayBuffer(total, n_env, alpha=0.6, beta=0.4)
    _, it = golden_stream(77, 30, n_env, 0.15, 0.05, True)
    stream = list(it)
    for ids, d in stream[:20]:
        if device_adds:
            buf.add_device(ts.Batch(**{k: torch.from_numpy(np.asarray(v)).cuda()
                                       for k, v in d.items() if k != "info"}), ids)
            buf.init_weight(torch.from_numpy(
                np.asarray(buf.last_index)[ids].astype(np.int64)).cuda())
        else:
            buf.add(ts.Batch(**d), buffer_ids=ids)
    buf.update_weight(np.array([0, 13, 25]), np.array([0.3, 1.7, 0.01], dtype=np.float32))
    blob = pickle.dumps(buf)
    new = pickle.loads(blob)
    _same_storage(new, buf)
    assert len(new) == len(buf)
    assert np.array_equal(new.last_index, buf.last_index)
    assert np.array_equal(new._lengths, buf._lengths)
    assert torch.equal(new.weight._value, buf.weight._value)
    q = np.arange(total)
    assert np.array_equal(new.prev(q), buf.prev(q)) and np.array_equal(new.next(q), buf.next(q))
    assert np.array_equal(new.unfinished_index(), buf.unfinished_index())
    assert torch.equal(new.sample_indices(0), buf.sample_indices(0))
    # children are views of the parent's bookkeeping again (manager.py:60-62)
    for k, child in enumerate(new.buffers):
        assert child._index == buf.buffers[k]._index and len(child) == len(buf.buffers[k])
    # both continue identically, on either add path
    for ids, d in stream[20:]:
        a = buf.add(ts.Batch(**d), buffer_ids=ids)
        b = new.add(ts.Batch(**d), buffer_ids=ids)
        for x, y in zip(a, b):
            assert np.array_equal(np.asarray(x), np.asarray(y))
    _same_storage(new, buf)
    u = torch.rand(64, dtype=torch.float64, device="cuda")
    assert torch.equal(new.sample_with_uniforms(u), buf.sample_with_uniforms(u))
