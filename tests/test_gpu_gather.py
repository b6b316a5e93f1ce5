"""Frame-stack gather on the GPU: ts_gather_stack through the device buffers
vs the reference's goldens, and through the raw C-ABI vs the CPU oracle.  Byte copies: bit-exact."""
import numpy as np
import pytest
import torch

from conftest import load_golden
from helpers import STACK_SPECS, random_rings, stack_stream
from oracle import oracle as orc

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("spec", STACK_SPECS, ids=[s[0] for s in STACK_SPECS])
def test_golden_stack_through_buffer(spec):
    """make_golden.gen_stack replayed through the device VectorReplayBuffer (Atari layout:
    stack_num, ignore_obs_next, save_only_last_obs; base.py:317-358, 374-378)."""
    import tianshou_fork_b200 as ts
    name, n_env, total, steps, k, ragged, fshape = spec
    c = load_golden("stack")[name]
    rng = np.random.default_rng(700 + STACK_SPECS.index(spec))
    buf = ts.VectorReplayBuffer(total, n_env, stack_num=k, ignore_obs_next=True,
                                save_only_last_obs=True)
    for ids, d in stack_stream(rng, steps, n_env, ragged, fshape, k):
        buf.add(ts.Batch(**d), buffer_ids=ids)
    assert np.array_equal(buf.obs.cpu().numpy(), c["frames"])
    assert np.array_equal(buf.done.cpu().numpy(), c["done"])
    q = c["query"]
    for index in (q, torch.from_numpy(q).cuda()):
        data = buf[index]
        assert np.array_equal(data.obs.cpu().numpy(), c["obs"])
        assert np.array_equal(data.obs_next.cpu().numpy(), c["obs_next"])
    assert np.array_equal(buf.get(q, "obs", stack_num=2).cpu().numpy(), c["get2"])
    # scalar index -> [k, ...] (axis = indices.ndim = 0)
    one = buf.get(int(q[5]), "obs")
    assert np.array_equal(one.cpu().numpy(), c["obs"][5])
    # 2-D index -> [a, b, k, ...]
    two = buf.get(q[:12].reshape(3, 4), "obs")
    assert np.array_equal(two.cpu().numpy(), c["obs"][:12].reshape((3, 4) + c["obs"].shape[1:]))
    with pytest.raises(IndexError):
        buf.get(np.array([total * 2]), "obs")


def test_golden_stack_single_ring():
    """ReplayBuffer(16, stack_num=3) with saved obs_next: both keys go through get()."""
    import tianshou_fork_b200 as ts
    c = load_golden("stack")["ring"]
    rng = np.random.default_rng(750)
    buf = ts.ReplayBuffer(16, stack_num=3)
    for _ in range(41):
        term = rng.random() < 0.2
        buf.add(ts.Batch(obs=rng.standard_normal(5).astype(np.float32), act=0,
                         rew=rng.standard_normal(), terminated=term,
                         truncated=(not term) and rng.random() < 0.1,
                         obs_next=rng.standard_normal(5).astype(np.float32), info={}))
    data = buf[c["query"]]
    assert np.array_equal(data.obs.cpu().numpy(), c["obs"])
    assert np.array_equal(data.obs_next.cpu().numpy(), c["obs_next"])


@pytest.mark.parametrize("row_shape,dtype,k,B,cap,M", [
    ((84, 84), np.uint8, 4, 16, 2048, 4096),    # Atari frames: 7056 B rows, bulk-copy path
    ((84, 84), np.uint8, 1, 16, 512, 300),      # stack of one
    ((27,), np.float32, 4, 8, 500, 1000),       # 108 B rows: 4-byte register path
    ((5,), np.uint8, 3, 3, 100, 257),           # odd rows: byte path
    ((64,), np.float32, 16, 4, 300, 513),       # deepest stack the kernel takes
])
def test_gather_stack_abi_vs_oracle(row_shape, dtype, k, B, cap, M):
    """ts_gather_stack through the raw C-ABI on random ring states, both outputs."""
    from tianshou_fork_b200 import _lib as L
    rng = np.random.default_rng(M + k)
    offset, done, last, lengths = random_rings(rng, B, cap, p_done=0.05)
    total = B * cap
    if dtype == np.uint8:
        frames = rng.integers(0, 256, (total,) + row_shape).astype(np.uint8)
    else:
        frames = rng.standard_normal((total,) + row_shape).astype(dtype)
    idx = rng.integers(0, total, M)
    idx[:B] = last                      # the newest slot of every ring (unfinished tails)
    idx[B:2 * B] = offset[:-1]          # the oldest possible slot
    want = orc.stack_gather(frames, idx, k, offset, done, last, lengths)
    want_next = orc.stack_gather(frames, idx, k, offset, done, last, lengths, from_next=True)
    d = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    g_frames, g_idx, g_off = d(frames), d(idx), d(offset)
    g_done, g_last, g_len = d(done.view(np.uint8)), d(last), d(lengths)
    out = torch.empty((M, k) + row_shape, dtype=g_frames.dtype, device="cuda")
    out_next = torch.empty_like(out)
    row_bytes = int(np.prod(row_shape)) * frames.itemsize
    for sub in (cap, 0):  # equal sub-buffers by division, then by binary search over offset
        out.fill_(0)
        out_next.fill_(0)
        L.check(L.lib().ts_gather_stack(L.ptr(g_frames), row_bytes, L.ptr(g_idx), M, k,
                                        L.ptr(g_off), B, sub, L.ptr(g_done), L.ptr(g_last),
                                        L.ptr(g_len), L.ptr(out), L.ptr(out_next), L.stream()),
                "ts_gather_stack")
        assert np.array_equal(out.cpu().numpy(), want)
        assert np.array_equal(out_next.cpu().numpy(), want_next)
    out.fill_(0)
    L.check(L.lib().ts_gather_stack(L.ptr(g_frames), row_bytes, L.ptr(g_idx), M, k, L.ptr(g_off),
                                    B, cap, L.ptr(g_done), L.ptr(g_last), L.ptr(g_len),
                                    L.ptr(out), None, L.stream()), "ts_gather_stack")
    assert np.array_equal(out.cpu().numpy(), want)


@pytest.mark.parametrize("m", [1, 33, 4096])
def test_fused_field_gather_equals_per_key_indexing(m):
    """buffer[indices]: act / rew / flags gathered by one ts_gather_fields launch
    (base.py:379-389 copies key by key) -- byte-exact against torch indexing, for 1-D and 2-D
    index arrays, negative indices and mixed row widths (u8 flags, f64 rew, f32[3] actions)."""
    import tianshou_fork_b200 as ts
    rng = np.random.default_rng(m)
    E, T = 4, 50
    buf = ts.VectorReplayBuffer(E * T, E)
    for t in range(T):
        buf.add(ts.Batch(obs=rng.standard_normal((E, 5)).astype(np.float32),
                         act=rng.standard_normal((E, 3)).astype(np.float32),
                         rew=rng.standard_normal(E), terminated=rng.random(E) < 0.1,
                         truncated=rng.random(E) < 0.05,
                         obs_next=rng.standard_normal((E, 5)).astype(np.float32), info={}))
    idx = rng.integers(-E * T, E * T, m)
    for index in (idx, torch.from_numpy(idx).cuda(), idx[: (m // 2) * 2].reshape(2, -1)):
        if np.size(index) == 0 if not isinstance(index, torch.Tensor) else index.numel() == 0:
            continue
        got = buf[index]
        ti = torch.as_tensor(index).cuda()
        for k in ("act", "rew", "terminated", "truncated", "done"):
            want = buf._meta[k][ti]
            assert got[k].dtype == want.dtype and got[k].shape == want.shape, k
            assert torch.equal(got[k], want), k
    lib_out = buf._take_fields(("act", "rew", "done"), torch.from_numpy(idx).cuda())
    assert torch.equal(lib_out[0], buf._meta.act[torch.from_numpy(idx).cuda()])
