"""Out-of-bounds write checks with guard bands (compute-sanitizer is not available on the GPU
pool): every output of the bulk-copy kernels is a window inside a larger sentinel-filled buffer;
the sentinels on both sides must survive the call, for sizes that are not tile multiples."""
import numpy as np
import pytest
import torch

from helpers import c2_rollout, random_rings

pytestmark = pytest.mark.gpu
PAD = 4096  # bytes on each side


class Guarded:
    """A tensor of `shape` / `dtype` carved out of a byte buffer with sentinel bands."""

    def __init__(self, shape, dtype):
        n = int(np.prod(shape)) * torch.empty(0, dtype=dtype).element_size()
        self.raw = torch.full((n + 2 * PAD,), 0xA5, dtype=torch.uint8, device="cuda")
        self.t = self.raw[PAD:PAD + n].view(dtype).reshape(shape)
        self.n = n

    def intact(self):
        r = self.raw
        return bool((r[:PAD] == 0xA5).all()) and bool((r[PAD + self.n:] == 0xA5).all())


@pytest.mark.parametrize("E,T", [(1025, 2048), (600, 4096 + 16), (9000, 120), (2100, 512),
                                 (1100, 1000), (33, 100)])
def test_gae_rows_writes_stay_inside(E, T):
    from tianshou_fork_b200 import _lib as L
    rew, v_s, v_next, term, trunc = c2_rollout(E, T, seed=1)
    d = lambda a: torch.from_numpy(a).cuda()
    g = [d(rew), d(v_s), d(v_next), d(term.view(np.uint8)), d(trunc.view(np.uint8))]
    adv, ret = Guarded((E, T), torch.float32), Guarded((E, T), torch.float32)
    L.check(L.lib().ts_gae_rows(*[L.ptr(t) for t in g], E, T, T, T, 0.99, 0.95, 0, 0, 0,
                                L.ptr(adv.t), L.ptr(ret.t), L.stream()), "ts_gae_rows")
    torch.cuda.synchronize()
    assert adv.intact() and ret.intact()
    assert bool(torch.isfinite(adv.t).all()) and bool(torch.isfinite(ret.t).all())


@pytest.mark.parametrize("N", [1 << 20, (1 << 20) + 7, 3_000_017, 2_500_000])
@pytest.mark.parametrize("with_idx", [False, True])
def test_gae_flat_writes_stay_inside(N, with_idx):
    from tianshou_fork_b200 import _lib as L
    rng = np.random.default_rng(N % 97)
    d = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    rew, v_s, v_n = (d(rng.standard_normal(N).astype(np.float32)) for _ in range(3))
    term = d((rng.random(N) < 1 / 300).view(np.uint8))
    trunc = d(np.zeros(N, np.uint8))
    idx = d(np.arange(N)) if with_idx else None
    unf = d(np.full(4, -1, np.int64)) if with_idx else None
    off = d(np.linspace(0, N, 5).astype(np.int64)) if with_idx else None
    adv, ret = Guarded((N,), torch.float32), Guarded((N,), torch.float32)
    nbytes = L.lib().ts_gae_flat_workspace_bytes(N)
    ws = Guarded((nbytes,), torch.uint8)
    ws.t.zero_()
    L.check(L.lib().ts_gae_flat(L.ptr(rew), L.ptr(v_s), L.ptr(v_n), L.ptr(term), L.ptr(trunc),
                                L.ptr(idx), L.ptr(term) if with_idx else None, L.ptr(unf),
                                L.ptr(off), 4, 0, N, 0.99, 0.95, 0, 0, 0, L.ptr(adv.t),
                                L.ptr(ret.t), L.ptr(ws.t), nbytes, L.stream()), "ts_gae_flat")
    torch.cuda.synchronize()
    assert adv.intact() and ret.intact() and ws.intact()
    assert bool(torch.isfinite(adv.t).all())


@pytest.mark.parametrize("row_shape,dtype,k,M", [((84, 84), torch.uint8, 4, 1001),
                                                 ((27,), torch.float32, 4, 333),
                                                 ((5,), torch.uint8, 3, 77)])
def test_gather_stack_writes_stay_inside(row_shape, dtype, k, M):
    from tianshou_fork_b200 import _lib as L
    rng = np.random.default_rng(5)
    B, cap = 8, 256
    offset, done, last, lengths = random_rings(rng, B, cap, p_done=0.05)
    total = B * cap
    frames = torch.randint(0, 200, (total,) + row_shape, device="cuda").to(dtype)
    d = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    idx = d(rng.integers(0, total, M))
    out, out_next = Guarded((M, k) + row_shape, dtype), Guarded((M, k) + row_shape, dtype)
    row_bytes = int(np.prod(row_shape)) * frames.element_size()
    g_off, g_done, g_last, g_len = d(offset), d(done.view(np.uint8)), d(last), d(lengths)
    L.check(L.lib().ts_gather_stack(L.ptr(frames), row_bytes, L.ptr(idx), M, k, L.ptr(g_off),
                                    B, cap, L.ptr(g_done), L.ptr(g_last), L.ptr(g_len),
                                    L.ptr(out.t), L.ptr(out_next.t), L.stream()),
            "ts_gather_stack")
    torch.cuda.synchronize()
    assert out.intact() and out_next.intact()


@pytest.mark.parametrize("size,bsz", [(1 << 20, 4096), (100_000, 1000), (5000, 33)])
def test_per_outputs_stay_inside(size, bsz):
    """Sample writes idx / weight windows; update writes only the tree (its own guard)."""
    from tianshou_fork_b200 import _lib as L
    rng = np.random.default_rng(3)
    bound = 1
    while bound < size:
        bound *= 2
    tree = Guarded((2 * bound,), torch.float64)
    tree.t.zero_()
    stamp = torch.zeros(size, dtype=torch.int32, device="cuda")
    counter = torch.zeros(4, dtype=torch.int32, device="cuda")
    err = torch.zeros(1, dtype=torch.int32, device="cuda")
    prio = torch.tensor([1.0, 1.0], dtype=torch.float64, device="cuda")
    all_idx = torch.arange(size, device="cuda")
    leaves = torch.from_numpy(rng.random(size) + 0.1).cuda()
    lib = L.lib()
    L.check(lib.ts_segtree_update(L.ptr(tree.t), size, bound, L.ptr(all_idx), L.ptr(leaves), 1, 1,
                                  size, L.ptr(stamp), L.ptr(counter), L.ptr(err), L.stream()),
            "update")
    u = torch.from_numpy(rng.random(bsz)).cuda()
    idx, w = Guarded((bsz,), torch.int64), Guarded((bsz,), torch.float64)
    L.check(lib.ts_per_sample(L.ptr(tree.t), bound, L.ptr(u), bsz, L.ptr(prio), 0.4, 1,
                              L.ptr(idx.t), L.ptr(w.t), None, L.ptr(err), L.stream()), "sample")
    td = torch.from_numpy(rng.standard_normal(bsz).astype(np.float32)).cuda()
    L.check(lib.ts_per_update(L.ptr(tree.t), size, bound, L.ptr(idx.t), L.ptr(td), 0, bsz, 0.6,
                              L.ptr(prio), L.ptr(stamp), L.ptr(counter), L.ptr(err), L.stream()),
            "per_update")
    torch.cuda.synchronize()
    assert idx.intact() and w.intact() and tree.intact()
    assert int(err.item()) == 0
    assert int(idx.t.min()) >= 0 and int(idx.t.max()) < size
