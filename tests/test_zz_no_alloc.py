"""ts_gae_flat_static (include/ts_b200.h): the flat GAE kernels with their scratch in a static
device array inside the library -- no workspace argument, no allocation.  Checked through the
raw C-ABI against ts_gae_flat (caller workspace) on the same inputs, two calls in a row, on both
kernels behind the entry point (pipelined ranges, decoupled look-back), plus the size limit.
Collected last on purpose (file name)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _inputs(N, seed, p_end):
    rng = np.random.default_rng(seed)
    rew = rng.standard_normal(N)
    v_s = rng.standard_normal(N).astype(np.float32)
    v_n = rng.standard_normal(N).astype(np.float32)
    term = rng.random(N) < p_end
    trunc = (rng.random(N) < p_end / 2) & ~term
    d = lambda a: torch.from_numpy(a).cuda()  # noqa: E731
    return [d(rew), d(v_s), d(v_n), d(term.view(np.uint8)), d(trunc.view(np.uint8))]


def _existing(L, g, N, gamma, lam, ws):
    adv = torch.empty(N, dtype=torch.float64, device="cuda")
    ret = torch.empty(N, dtype=torch.float64, device="cuda")
    L.check(L.lib().ts_gae_flat(*[L.ptr(t) for t in g], None, None, None, None, 0, 0, N, gamma,
                                lam, 1, 0, 1, L.ptr(adv), L.ptr(ret), L.ptr(ws), ws.numel(),
                                L.stream()), "ts_gae_flat")
    return adv, ret


def _static(L, g, N, gamma, lam):
    adv = torch.full((N,), float("nan"), dtype=torch.float64, device="cuda")
    ret = torch.full((N,), float("nan"), dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    L.check(L.lib().ts_gae_flat_static(*[L.ptr(t) for t in g], None, None, None, None, 0, 0, N,
                                       gamma, lam, 1, 0, 1, L.ptr(adv), L.ptr(ret), L.stream()),
            "ts_gae_flat_static")
    assert torch.cuda.memory_allocated() == before
    return adv, ret


@pytest.mark.parametrize("N,gamma,lam,p_end", [
    (3_000_017, 0.99, 0.95, 0.002),   # pipelined ranges, ragged tail
    (1 << 21, 1.0, 1.0, 0.0),         # no episode end, no discount: neighbour flags in the scratch
    (300_001, 0.99, 0.95, 0.0005),    # decoupled look-back (ticket + status cleared per call)
    (777, 0.9, 0.8, 0.05),
])
def test_static_scratch_entry_matches_workspace_entry(N, gamma, lam, p_end):
    from tianshou_fork_b200 import _lib as L
    lib = L.lib()
    assert N <= lib.ts_gae_flat_static_max_n()
    ws = torch.zeros(lib.ts_gae_flat_workspace_bytes(N), dtype=torch.uint8, device="cuda")
    for call in range(2):  # two calls in a row on the shared scratch, different data
        g = _inputs(N, seed=1000 * call + N % 997, p_end=p_end)
        want_adv, want_ret = _existing(L, g, N, gamma, lam, ws)
        got_adv, got_ret = _static(L, g, N, gamma, lam)
        a, b = got_adv.cpu().numpy(), want_adv.cpu().numpy()
        scale = np.abs(b).max()
        # the tolerance of tests/test_gpu_gae.py for f64 outputs (1e-9); same kernels, same
        # composition order, so the two are in fact expected to be bit-equal
        assert np.all(np.abs(a - b) <= 1e-9 * np.abs(b) + 1e-11 * scale), (N, call)
        r, q = got_ret.cpu().numpy(), want_ret.cpu().numpy()
        assert np.all(np.abs(r - q) <= 1e-9 * np.abs(q) + 1e-11 * np.abs(q).max()), (N, call)


def test_static_scratch_entry_rejects_sizes_beyond_its_capacity():
    from tianshou_fork_b200 import _lib as L
    lib = L.lib()
    cap = lib.ts_gae_flat_static_max_n()
    assert cap == 1 << 24
    one = torch.zeros(16, dtype=torch.float64, device="cuda")
    flags = torch.zeros(16, dtype=torch.uint8, device="cuda")
    code = lib.ts_gae_flat_static(L.ptr(one), L.ptr(one), L.ptr(one), L.ptr(flags), L.ptr(flags),
                                  None, None, None, None, 0, 0, cap + 1, 0.99, 0.95, 1, 1, 1,
                                  L.ptr(one), L.ptr(one), L.stream())
    assert code == -5  # TS_ERR_TOO_LARGE, before any launch
    assert b"capacity" in lib.ts_error_string(code)
    with pytest.raises(L.TsError, match="capacity"):
        L.check(code, "ts_gae_flat_static")
