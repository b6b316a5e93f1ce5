"""Device-side add: ts_add_step through ReplayBufferManager.add_device vs the
reference's bookkeeping goldens (make_golden.gen_index) and vs the host add() path.  Integer /
byte work: bit-exact."""
import numpy as np
import pytest
import torch

from conftest import load_golden
from helpers import INDEX_SPECS, golden_stream

pytestmark = pytest.mark.gpu


def _dev_batch(ts, d):
    return ts.Batch(**{k: (torch.from_numpy(np.asarray(v)).cuda() if k != "info" else v)
                       for k, v in d.items() if k != "info"})


@pytest.mark.parametrize("spec", INDEX_SPECS, ids=[s[0] for s in INDEX_SPECS])
def test_add_device_matches_goldens_and_host_add(spec):
    import tianshou_fork_b200 as ts
    name, n_env, total, steps, ragged = spec
    c = load_golden("index")[name]
    seed = 200 + INDEX_SPECS.index(spec)
    dev_buf = ts.VectorReplayBuffer(total, n_env)
    host_buf = ts.VectorReplayBuffer(total, n_env)
    _, it = golden_stream(seed, steps, n_env, 0.15, 0.05, ragged)
    for step, (ids, d) in enumerate(it):
        hp, her, hel, hei = host_buf.add(ts.Batch(**d), buffer_ids=ids)
        full = ids.size == n_env and step % 2 == 0
        dp, der, del_, dei = dev_buf.add_device(_dev_batch(ts, d), None if full else ids)
        assert np.array_equal(dp.cpu().numpy(), hp)
        assert np.array_equal(der.cpu().numpy(), her)
        assert np.array_equal(del_.cpu().numpy(), hel)
        assert np.array_equal(dei.cpu().numpy(), hei)
    # the reference's bookkeeping (test_buffer.py:710-712, 836-841 style state)
    assert np.array_equal(dev_buf.done.cpu().numpy(), c["done"])
    assert np.array_equal(dev_buf.last_index, c["last_index"])
    assert np.array_equal(dev_buf._lengths, c["lengths"])
    assert np.array_equal(np.array([b._index for b in dev_buf.buffers]), c["ring_index"])
    assert np.array_equal(dev_buf.unfinished_index(), c["unfinished"])
    assert np.array_equal(dev_buf.prev(c["query"]), c["prev"])
    assert np.array_equal(dev_buf.next(c["query"]), c["next"])
    assert np.array_equal(dev_buf.sample_indices(0).cpu().numpy(), c["all_indices"])
    # device copies of the state the kernel advanced in place
    st = dev_buf._dev_add
    assert np.array_equal(st["last_index"].cpu().numpy(), c["last_index"])
    assert np.array_equal(st["lengths"].cpu().numpy(), c["lengths"])
    assert np.array_equal(st["ring"].cpu().numpy()[:, 0], c["ring_index"])
    assert int(st["err"].item()) == 0
    # every stored key, byte for byte
    for k in ("obs", "act", "rew", "terminated", "truncated", "done", "obs_next"):
        assert torch.equal(dev_buf._meta[k], host_buf._meta[k]), k
    assert len(dev_buf) == len(host_buf)


def test_add_device_interleaved_with_host_add():
    """Episode statistics move between host and device copies when the two paths alternate."""
    import tianshou_fork_b200 as ts
    n_env, total = 5, 60
    a, b = ts.VectorReplayBuffer(total, n_env), ts.VectorReplayBuffer(total, n_env)
    _, it = golden_stream(31, 40, n_env, 0.2, 0.05, True)
    for step, (ids, d) in enumerate(it):
        want = a.add(ts.Batch(**d), buffer_ids=ids)
        if (step // 3) % 2 == 0:
            got = [t.cpu().numpy() for t in b.add_device(_dev_batch(ts, d), ids)]
        else:
            got = b.add(ts.Batch(**d), buffer_ids=ids)
        for g, w in zip(got, want):
            assert np.array_equal(np.asarray(g), np.asarray(w)), step
    for k in ("rew", "done", "obs"):
        assert torch.equal(a._meta[k], b._meta[k])
    assert np.array_equal(a.last_index, b.last_index)


def test_add_device_atari_frames_and_stack():
    """save_only_last_obs + ignore_obs_next storage filled on the device, read back stacked."""
    import tianshou_fork_b200 as ts
    from helpers import STACK_SPECS, stack_stream
    name, n_env, total, steps, k, ragged, fshape = STACK_SPECS[0]
    c = load_golden("stack")[name]
    rng = np.random.default_rng(700)
    buf = ts.VectorReplayBuffer(total, n_env, stack_num=k, ignore_obs_next=True,
                                save_only_last_obs=True)
    for ids, d in stack_stream(rng, steps, n_env, ragged, fshape, k):
        buf.add_device(_dev_batch(ts, d), ids)
    assert np.array_equal(buf.obs.cpu().numpy(), c["frames"])
    data = buf[c["query"]]
    assert np.array_equal(data.obs.cpu().numpy(), c["obs"])
    assert np.array_equal(data.obs_next.cpu().numpy(), c["obs_next"])


@pytest.mark.parametrize("n_env,frame", [(16, (27,)), (10, (84, 84)), (64, (3,))])
def test_packed_host_add_equals_per_key_assignment(n_env, frame):
    """ReplayBufferManager.add with NumPy steps: one packed upload + ts_scatter_fields
    (manager.py:160 assigns key by key) -- storage byte-equal to the per-key path, over ragged
    ids, ring wrap-around and u8 frames."""
    import tianshou_fork_b200 as ts
    dtype = np.uint8 if len(frame) == 2 else np.float32
    bufs = [ts.VectorReplayBuffer(n_env * 12, n_env) for _ in range(2)]
    bufs[1]._store_packed = lambda ptrs, batch: False  # reference-style per-key assignment
    rng = np.random.default_rng(2)
    for t in range(40):
        ids = np.arange(n_env)
        if t % 3 == 1:
            ids = ids[rng.random(n_env) < 0.6]
            if ids.size == 0:
                continue
        m = ids.size
        step = dict(obs=(rng.random((m,) + frame) * 255).astype(dtype),
                    act=rng.integers(0, 5, m), rew=rng.standard_normal(m),
                    terminated=rng.random(m) < 0.1, truncated=rng.random(m) < 0.05,
                    obs_next=(rng.random((m,) + frame) * 255).astype(dtype), info={})
        outs = [b.add(ts.Batch(**step), buffer_ids=ids) for b in bufs]
        for x, y in zip(*outs):
            assert np.array_equal(x, y)
    for k in ("obs", "act", "rew", "terminated", "truncated", "done", "obs_next"):
        assert torch.equal(bufs[0]._meta[k], bufs[1]._meta[k]), k


def test_add_device_on_a_prioritized_buffer_initialises_the_leaves():
    """``add_device`` on a PrioritizedVectorReplayBuffer calls ``init_weight`` on the slots it
    wrote like ``add`` does (prio.py:50-57): same tree as a twin filled through the host path,
    and transitions added on the device are sampled."""
    import tianshou_fork_b200 as ts
    rng = np.random.default_rng(4)
    E, cap = 4, 16
    dev_buf = ts.PrioritizedVectorReplayBuffer(E * cap, E, alpha=0.7, beta=0.5)
    host_buf = ts.PrioritizedVectorReplayBuffer(E * cap, E, alpha=0.7, beta=0.5)
    for t in range(23):  # wraps
        ids = np.arange(E) if t % 3 else np.array([0, 2])
        m = ids.size
        step = dict(obs=rng.standard_normal((m, 3)).astype(np.float32),
                    act=rng.integers(0, 4, m), rew=rng.standard_normal(m),
                    terminated=rng.random(m) < 0.2, truncated=np.zeros(m, bool))
        host_buf.add(ts.Batch(info={}, **step), buffer_ids=ids)
        dev_buf.add_device(ts.Batch(**{k: torch.from_numpy(v).cuda() for k, v in step.items()}),
                           buffer_ids=ids)
        if t == 9:  # raise max_prio so later init_weight values differ from 1
            for b in (host_buf, dev_buf):
                b.update_weight(np.array([1, 17]), np.array([3.0, 0.25]))
    assert torch.equal(dev_buf.weight._value, host_buf.weight._value)
    assert dev_buf.weight.reduce() > 0
    np.random.seed(3)
    a = dev_buf.sample_indices(64).cpu().numpy()
    np.random.seed(3)
    b = host_buf.sample_indices(64).cpu().numpy()
    assert np.array_equal(a, b)
    with pytest.raises(NotImplementedError):
        ts.PrioritizedReplayBuffer(8, 0.6, 0.4).add_device(ts.Batch())
