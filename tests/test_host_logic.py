"""CPU tests of host-side logic that needs no kernel: Batch container and split boundaries."""
import numpy as np
import torch

from tianshou_fork_b200.data.batch import Batch, create_value
from tianshou_fork_b200.policy.ppo import _split_bounds


def test_batch_index_assign_len():
    b = Batch(a=np.arange(6), obs=Batch(x=np.arange(12).reshape(6, 2)), info={})
    assert len(b) == 6 and "a" in b and b.info.is_empty()
    sub = b[np.array([4, 1])]
    assert np.array_equal(sub.a, [4, 1]) and np.array_equal(sub.obs.x, [[8, 9], [2, 3]])
    store = create_value(b, 10, stack=False)
    assert store.a.shape == (10,) and store.obs.x.shape == (10, 2)
    store[np.array([7, 8])] = sub
    assert np.array_equal(store.a[7:9], [4, 1])
    t = Batch(a=torch.arange(6))
    assert torch.equal(t[np.array([0, 5])].a, torch.tensor([0, 5]))
    assert b.pop("a") is not None and "a" not in b


def test_batch_split_matches_reference_boundaries():
    """Batch.split (tianshou/data/batch.py:896-912): merge_last folds the short tail."""
    b = Batch(a=np.arange(10))
    sizes = [len(m) for m in b.split(4, shuffle=False, merge_last=True)]
    assert sizes == [4, 6]
    sizes = [len(m) for m in b.split(4, shuffle=False, merge_last=False)]
    assert sizes == [4, 4, 2]
    sizes = [len(m) for m in b.split(5, shuffle=False, merge_last=True)]
    assert sizes == [5, 5]
    assert _split_bounds(10, 4, True) == [(0, 4), (4, 10)]
    assert _split_bounds(10, 4, False) == [(0, 4), (4, 8), (8, 10)]
    assert _split_bounds(10, -1, True) == [(0, 10)]
    np.random.seed(3)
    got = np.concatenate([m.a for m in b.split(3)])
    np.random.seed(3)
    assert np.array_equal(got, np.random.permutation(10))


def test_batch_conveniences():
    """Batch.update / copy / shape / cat / arithmetic on host leaves (no GPU needed)."""
    import numpy as np
    from tianshou_fork_b200.data.batch import Batch
    a = Batch(obs=np.arange(6.0).reshape(3, 2), info=Batch(t=np.arange(3)), empty=Batch())
    b = a.copy()
    b.obs[0, 0] = 99.0
    assert a.obs[0, 0] == 0.0 and a.shape == [3]
    c = Batch.cat([a, b])
    assert len(c) == 6 and c.info.t.tolist() == [0, 1, 2, 0, 1, 2] and c.empty.is_empty()
    a.cat_(b)
    assert len(a) == 6 and a.obs[3, 0] == 99.0
    a.update(rew=np.ones(6), obs=np.zeros((6, 2)))
    assert set(a.keys()) == {"obs", "info", "empty", "rew"} and a.obs.sum() == 0
    d = Batch(x=np.array([1.0, 2.0]))
    e = d * 2 + 1
    assert e.x.tolist() == [3.0, 5.0] and d.x.tolist() == [1.0, 2.0]
    d /= 2
    assert d.x.tolist() == [0.5, 1.0]
    parts = list(Batch(x=np.arange(10)).split(4, shuffle=False, merge_last=True))
    assert [len(p) for p in parts] == [4, 6]


def test_converters():
    import numpy as np
    import torch
    from tianshou_fork_b200.data import Batch, to_numpy, to_torch, to_torch_as
    b = Batch(a=torch.arange(3), n=Batch(x=np.ones(2)))
    nb = to_numpy(b)
    assert isinstance(nb.a, np.ndarray) and isinstance(b.a, torch.Tensor)  # a copy is converted
    t = to_torch(nb, dtype=torch.float32)
    assert t.a.dtype == torch.int64 and t.n.x.dtype == torch.float32  # dtype applies to floats
    y = torch.zeros(1, dtype=torch.float64)
    assert to_torch_as(np.ones(2, np.float32), y).dtype == torch.float64
    assert to_numpy(3).shape == () and to_numpy([1, 2]).tolist() == [1, 2]
