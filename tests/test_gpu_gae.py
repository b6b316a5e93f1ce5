"""GAE parity on the GPU: CUDA path (through the C-ABI) vs the CPU oracle and the golden
fixtures produced by the reference.  Tolerance: 1e-5 relative (BASELINE.json north_star);
the f64-output path is additionally held to 1e-9."""
import numpy as np
import pytest
import torch

from conftest import load_golden
from helpers import GAE_SPECS, c2_rollout, fill
from oracle import oracle as orc

pytestmark = pytest.mark.gpu

RTOL = 1e-5  # north_star: "within 1e-5 relative for trajectories up to 4096 steps"


def _close(a, b, rtol=RTOL, scale=None):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    scale = np.abs(b).max() if scale is None else scale
    # relative to each element, with an absolute floor tied to the array's magnitude: an
    # advantage that cancels to ~0 cannot carry 1e-5 relative accuracy from f32 inputs
    return np.all(np.abs(a - b) <= rtol * np.abs(b) + rtol * 1e-2 * scale)


@pytest.mark.parametrize("spec", GAE_SPECS, ids=[s[0] for s in GAE_SPECS])
def test_golden_gae_through_buffer(spec):
    """make_golden.gen_gae replayed through the device VectorReplayBuffer + drop-in API."""
    import tianshou_fork_b200 as ts
    name, n_env, steps, cap, gamma, lam, pt, ptr, ragged = spec
    c = load_golden("gae")[name]
    buf = ts.VectorReplayBuffer(cap * n_env, n_env)
    fill(buf, ts.Batch, 100 + GAE_SPECS.index(spec), steps, n_env, pt, ptr, ragged)
    # bookkeeping and stored flags are integer/byte work: bit-exact
    assert np.array_equal(buf.done.cpu().numpy(), c["buf_done"])
    assert np.array_equal(buf.last_index, c["last_index"])
    assert np.array_equal(buf._lengths, c["lengths"])
    assert np.array_equal(buf.unfinished_index(), c["unfinished"])
    batch, indices = buf.sample(0)
    assert np.array_equal(indices.cpu().numpy(), c["indices"])
    assert np.array_equal(batch.rew.cpu().numpy(), c["buf_rew"][c["indices"]])
    # NumPy in -> NumPy f64 out, like the reference
    ret, adv = ts.BasePolicy.compute_episodic_return(
        batch, buf, indices, v_s_=c["v_s_"].copy(), v_s=c["v_s"].copy(), gamma=gamma,
        gae_lambda=lam)
    assert isinstance(ret, np.ndarray) and ret.dtype == np.float64
    assert _close(adv, c["adv"], 1e-9) and _close(ret, c["returns"], 1e-9), name
    # tensor in -> f32 tensor out, no host round trip
    dev = buf.device
    ret32, adv32 = ts.BasePolicy.compute_episodic_return(
        batch, buf, indices, v_s_=torch.from_numpy(c["v_s_"]).to(dev),
        v_s=torch.from_numpy(c["v_s"]).to(dev), gamma=gamma, gae_lambda=lam)
    assert ret32.is_cuda and ret32.dtype == torch.float32
    assert _close(adv32.cpu().numpy(), c["adv"]) and _close(ret32.cpu().numpy(), c["returns"])
    if name == "small":
        r2, a2 = ts.BasePolicy.compute_episodic_return(
            batch, buf, indices, v_s_=c["v_s_"].copy(), v_s=None, gamma=gamma, gae_lambda=lam)
        assert _close(r2, c["returns_roll"], 1e-9) and _close(a2, c["adv_roll"], 1e-9)
        r3, a3 = ts.BasePolicy.compute_episodic_return(
            ts.Batch(rew=batch.rew.cpu().numpy(), terminated=batch.terminated.cpu().numpy(),
                     truncated=batch.truncated.cpu().numpy()),
            buf, indices.cpu().numpy(), v_s_=None, v_s=None, gamma=gamma, gae_lambda=1.0)
        assert _close(r3, c["returns_mc"], 1e-9) and _close(a3, c["adv_mc"], 1e-9)


def test_canonical_rows_path_is_taken_and_matches_flat():
    """sample(0) of an un-wrapped buffer goes through ts_gae_rows; same numbers as ts_gae_flat."""
    import tianshou_fork_b200 as ts
    name, n_env, steps, cap, gamma, lam, pt, ptr, ragged = GAE_SPECS[3]  # long4096
    c = load_golden("gae")[name]
    buf = ts.VectorReplayBuffer(cap * n_env, n_env)
    fill(buf, ts.Batch, 103, steps, n_env, pt, ptr, ragged)
    batch, indices = buf.sample(0)
    assert buf.is_canonical(indices) == (n_env, steps)
    dev = buf.device
    vs, vn = torch.from_numpy(c["v_s"]).to(dev), torch.from_numpy(c["v_s_"]).to(dev)
    r_rows, a_rows = ts.BasePolicy.compute_episodic_return(
        batch, buf, indices, vn, vs, gamma, lam, out_dtype=torch.float64)
    r_flat, a_flat = ts.BasePolicy.compute_episodic_return(
        batch, buf, indices.clone(), vn, vs, gamma, lam, out_dtype=torch.float64)
    assert buf.is_canonical(indices.clone()) is None
    assert _close(a_rows.cpu().numpy(), c["adv"], 1e-9)
    assert _close(a_flat.cpu().numpy(), c["adv"], 1e-9)
    assert _close(r_rows.cpu().numpy(), r_flat.cpu().numpy(), 1e-12)


@pytest.mark.parametrize("E,T", [(1, 1), (3, 7), (5, 8), (4, 33), (7, 255), (2, 256), (3, 257),
                                 (6, 1000), (2, 2048), (3, 2049), (2, 4096), (1, 10000),
                                 (64, 2048), (300, 128)])
@pytest.mark.parametrize("dt", ["f32", "f64"])
def test_rows_kernel_vs_oracle(E, T, dt):
    """ts_gae_rows through the raw C-ABI on ragged / unaligned / multi-chunk shapes."""
    from tianshou_fork_b200 import _lib as L
    rew, v_s, v_next, term, trunc = c2_rollout(E, T, seed=E * 131 + T, p_term=1 / 50,
                                               p_trunc=1 / 100)
    ref_ret, ref_adv = orc.episodic_return_rows(rew, v_s, v_next, term, trunc, 0.99, 0.95)
    tdt = torch.float32 if dt == "f32" else torch.float64
    d = lambda a, t=None: torch.from_numpy(a).cuda().to(t) if t else torch.from_numpy(a).cuda()
    g_rew, g_vs, g_vn = d(rew, tdt), d(v_s, tdt), d(v_next, tdt)
    g_term, g_trunc = d(term.view(np.uint8)), d(trunc.view(np.uint8))
    adv = torch.empty(E, T, dtype=tdt, device="cuda")
    ret = torch.empty(E, T, dtype=tdt, device="cuda")
    code = L.dtype_code(g_rew)
    L.check(L.lib().ts_gae_rows(L.ptr(g_rew), L.ptr(g_vs), L.ptr(g_vn), L.ptr(g_term),
                                L.ptr(g_trunc), E, T, T, T, 0.99, 0.95, code, code, code,
                                L.ptr(adv), L.ptr(ret), L.stream()), "ts_gae_rows")
    tol = RTOL if dt == "f32" else 1e-9
    assert _close(adv.cpu().numpy(), ref_adv, tol)
    assert _close(ret.cpu().numpy(), ref_ret, tol)


def test_rows_kernel_strided_buffer_rows():
    """rew/flags rows `buf_stride` apart (sub-buffer capacity), values `val_stride` apart."""
    from tianshou_fork_b200 import _lib as L
    E, T, cap = 9, 100, 160
    rew, v_s, v_next, term, trunc = c2_rollout(E, T, seed=5, p_term=1 / 20)
    ref_ret, ref_adv = orc.episodic_return_rows(rew, v_s, v_next, term, trunc, 0.97, 0.9)
    pad = lambda a: np.concatenate([a, np.full((E, cap - T), 7, a.dtype)], axis=1)
    g_rew = torch.from_numpy(pad(rew)).cuda()
    g_term = torch.from_numpy(pad(term.view(np.uint8))).cuda()
    g_trunc = torch.from_numpy(pad(trunc.view(np.uint8))).cuda()
    g_vs, g_vn = torch.from_numpy(v_s).cuda(), torch.from_numpy(v_next).cuda()
    adv = torch.empty(E, T, device="cuda")
    ret = torch.empty(E, T, device="cuda")
    L.check(L.lib().ts_gae_rows(L.ptr(g_rew), L.ptr(g_vs), L.ptr(g_vn), L.ptr(g_term),
                                L.ptr(g_trunc), E, T, cap, T, 0.97, 0.9, 0, 0, 0, L.ptr(adv),
                                L.ptr(ret), L.stream()), "ts_gae_rows")
    assert _close(adv.cpu().numpy(), ref_adv) and _close(ret.cpu().numpy(), ref_ret)


@pytest.mark.parametrize("variant", [1, 2])
@pytest.mark.parametrize("N,p_end", [(1, 0.5), (15, 0.2), (16, 0.0), (2047, 0.01), (2048, 0.0),
                                     (2049, 0.001), (100_000, 0.0), (300_001, 0.0005),
                                     (1_000_000, 0.005), (3_000_017, 0.0)])
def test_flat_kernel_lookback_vs_oracle(N, p_end, variant):
    """ts_gae_flat, both kernels (1: decoupled look-back per tile, 2: pipelined ranges with
    read-ahead carry): one long sequence; p_end = 0 leaves no episode boundary, so the carry
    has to come from the discount decay (gamma*lambda = 0.94) or from the neighbour chain."""
    from tianshou_fork_b200 import _lib as L
    L.check(L.lib().ts_tune_set(b"gae_flat_variant", variant), "tune")
    rng = np.random.default_rng(N)
    rew = rng.standard_normal(N)
    v_s = rng.standard_normal(N).astype(np.float32)
    v_n = rng.standard_normal(N).astype(np.float32)
    term = rng.random(N) < p_end
    trunc = (rng.random(N) < p_end / 2) & ~term
    end = term | trunc
    ref_adv = orc.gae_return(v_s, v_n * ~term, rew, end, 0.99, 0.95)
    d = lambda a: torch.from_numpy(a).cuda()
    g = [d(rew), d(v_s), d(v_n), d(term.view(np.uint8)), d(trunc.view(np.uint8))]
    adv = torch.empty(N, dtype=torch.float64, device="cuda")
    ret = torch.empty(N, dtype=torch.float64, device="cuda")
    nbytes = L.lib().ts_gae_flat_workspace_bytes(N)
    ws = torch.zeros(nbytes, dtype=torch.uint8, device="cuda")
    for _ in range(3):  # workspace reuse across calls
        L.check(L.lib().ts_gae_flat(*[L.ptr(t) for t in g], None, None, None, None, 0, 0, N,
                                    0.99, 0.95, 1, 0, 1, L.ptr(adv), L.ptr(ret), L.ptr(ws),
                                    nbytes, L.stream()), "ts_gae_flat")
    L.check(L.lib().ts_tune_set(b"gae_flat_variant", 0), "tune")
    assert _close(adv.cpu().numpy(), ref_adv, 1e-9)
    assert _close(ret.cpu().numpy(), ref_adv + v_s, 1e-9)


@pytest.mark.parametrize("gl", [(1.0, 1.0), (0.999, 1.0)])
def test_flat_kernel_neighbour_chain(gl):
    """No episode end and (almost) no discount: read-ahead cannot resolve the carry, every
    range takes its neighbour's published head value (the Monte-Carlo return of pg.py:108-112
    over one 600 000-step sequence)."""
    from tianshou_fork_b200 import _lib as L
    gamma, lam = gl
    N = 600_000
    rng = np.random.default_rng(3)
    rew = rng.standard_normal(N)
    v_s = rng.standard_normal(N).astype(np.float32)
    v_n = rng.standard_normal(N).astype(np.float32)
    z = np.zeros(N, dtype=bool)
    ref_adv = orc.gae_return(v_s, v_n, rew, z, gamma, lam)
    d = lambda a: torch.from_numpy(a).cuda()
    g = [d(rew), d(v_s), d(v_n), d(z.view(np.uint8)), d(z.view(np.uint8))]
    adv = torch.empty(N, dtype=torch.float64, device="cuda")
    ret = torch.empty(N, dtype=torch.float64, device="cuda")
    nbytes = L.lib().ts_gae_flat_workspace_bytes(N)
    ws = torch.zeros(nbytes, dtype=torch.uint8, device="cuda")
    def call():
        L.check(L.lib().ts_gae_flat(*[L.ptr(t) for t in g], None, None, None, None, 0, 0, N,
                                    gamma, lam, 1, 0, 1, L.ptr(adv), L.ptr(ret), L.ptr(ws),
                                    nbytes, L.stream()), "ts_gae_flat")

    call()
    assert _close(adv.cpu().numpy(), ref_adv, 1e-9)
    # The same workspace again with other rewards, then through the look-back kernel and back:
    # head values of an earlier call must never be taken for this call's (flags carry a call tag).
    for variant in (2, 2, 1, 2):
        rew2 = rng.standard_normal(N)
        g[0] = d(rew2)
        L.check(L.lib().ts_tune_set(b"gae_flat_variant", variant), "tune")
        adv.fill_(float("nan"))
        call()
        L.check(L.lib().ts_tune_set(b"gae_flat_variant", 0), "tune")
        assert _close(adv.cpu().numpy(), orc.gae_return(v_s, v_n, rew2, z, gamma, lam), 1e-9)


@pytest.mark.parametrize("B,cap,rotate", [(37, 1000, True), (700, 5, False), (300, 13, False),
                                          (300, 13, True)])
@pytest.mark.parametrize("variant", [1, 2])
def test_flat_kernel_with_indices_vs_oracle(variant, B, cap, rotate):
    """Gathered form through the raw C-ABI: value mask from buffer.terminated[indices], end
    flags from the unfinished index of each sub-buffer (policy/base.py:335, 379-380)."""
    from tianshou_fork_b200 import _lib as L
    L.check(L.lib().ts_tune_set(b"gae_flat_variant", variant), "tune")
    rng = np.random.default_rng(11)
    # a batch that visits every sub-buffer's slots in a rotated order (wrapped rings); without
    # rotation 8-step strips are consecutive and straddle one or several short sub-buffers
    rot = rng.integers(0, cap, B) if rotate else np.zeros(B, dtype=np.int64)
    indices = np.concatenate([k * cap + (rot[k] + np.arange(cap)) % cap for k in range(B)])
    N = indices.size
    buf_term = rng.random(B * cap) < 1 / 80
    buf_trunc = (rng.random(B * cap) < 1 / 300) & ~buf_term
    unfinished = np.array([k * cap + (rot[k] - 1) % cap for k in range(B)])  # newest slot per ring
    unfinished[::5] = -1
    rew = rng.standard_normal(N)
    v_s = rng.standard_normal(N).astype(np.float32)
    v_n = rng.standard_normal(N).astype(np.float32)
    term, trunc = buf_term[indices], buf_trunc[indices]
    end = term | trunc | np.isin(indices, unfinished)
    ref_adv = orc.gae_return(v_s, v_n * ~term, rew, end, 0.99, 0.95)
    d = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    offset = d(np.arange(B + 1) * cap)
    adv = torch.empty(N, dtype=torch.float32, device="cuda")
    ret = torch.empty(N, dtype=torch.float32, device="cuda")
    nbytes = L.lib().ts_gae_flat_workspace_bytes(N)
    ws = torch.zeros(nbytes, dtype=torch.uint8, device="cuda")
    g = [d(rew), d(v_s), d(v_n), d(term.view(np.uint8)), d(trunc.view(np.uint8)), d(indices),
         d(buf_term.view(np.uint8)), d(unfinished), offset]
    for sub in (cap, 0):  # equal sub-buffers by division, then by binary search over offset
        L.check(L.lib().ts_gae_flat(*[L.ptr(t) for t in g], B, sub, N, 0.99, 0.95, 1, 0, 0,
                                    L.ptr(adv), L.ptr(ret), L.ptr(ws), nbytes, L.stream()),
                "ts_gae_flat")
        assert _close(adv.cpu().numpy(), ref_adv) and _close(ret.cpu().numpy(), ref_adv + v_s)
    L.check(L.lib().ts_tune_set(b"gae_flat_variant", 0), "tune")


def test_c2_full_size_properties():
    """BASELINE.json configs[1] at full size (4096 x 2048): the WHOLE array against the oracle
    (SURVEY.md 8d), plus size-independent properties."""
    from tianshou_fork_b200 import _lib as L
    E, T = 4096, 2048
    rew, v_s, v_next, term, trunc = c2_rollout(E, T, seed=0)
    d = lambda a: torch.from_numpy(a).cuda()
    g_rew, g_vs, g_vn = d(rew), d(v_s), d(v_next)
    g_term, g_trunc = d(term.view(np.uint8)), d(trunc.view(np.uint8))
    adv = torch.empty(E, T, device="cuda")
    ret = torch.empty(E, T, device="cuda")

    def run(r, a_out, r_out):
        L.check(L.lib().ts_gae_rows(L.ptr(r), L.ptr(g_vs), L.ptr(g_vn), L.ptr(g_term),
                                    L.ptr(g_trunc), E, T, T, T, 0.99, 0.95, 0, 0, 0,
                                    L.ptr(a_out), L.ptr(r_out), L.stream()), "ts_gae_rows")

    run(g_rew, adv, ret)
    a, r = adv.cpu().numpy(), ret.cpu().numpy()
    ref_ret, ref_adv = orc.episodic_return_rows(rew, v_s, v_next, term, trunc, 0.99, 0.95,
                                                threads=orc.max_threads())
    assert _close(a, ref_adv) and _close(r, ref_ret)  # all 8 388 608 transitions
    # f64 outputs of the same launch shape: 1e-9
    adv64 = torch.empty(E, T, dtype=torch.float64, device="cuda")
    ret64 = torch.empty(E, T, dtype=torch.float64, device="cuda")
    L.check(L.lib().ts_gae_rows(L.ptr(g_rew), L.ptr(g_vs), L.ptr(g_vn), L.ptr(g_term),
                                L.ptr(g_trunc), E, T, T, T, 0.99, 0.95, 0, 0, 1,
                                L.ptr(adv64), L.ptr(ret64), L.stream()), "ts_gae_rows")
    assert _close(adv64.cpu().numpy(), ref_adv, 1e-9) and _close(ret64.cpu().numpy(), ref_ret, 1e-9)
    del adv64, ret64
    # returns = advantage + v_s  (policy/base.py:382)
    assert np.allclose(r, a + v_s, rtol=0, atol=1e-5 * np.abs(r).max())
    # one-step Bellman identity away from episode ends: A_t = delta_t + gl * A_{t+1}
    end = term | trunc
    end[:, -1] = True
    delta = rew.astype(np.float64) + 0.99 * v_next * ~term - v_s
    rhs = delta.copy()
    rhs[:, :-1] += 0.99 * 0.95 * (~end[:, :-1]) * a[:, 1:]
    assert np.allclose(a, rhs, rtol=0, atol=2e-5 * np.abs(a).max())
    # linearity in the reward: A(2*rew) - A(rew) == A(rew) - A(0)
    adv2, adv0 = torch.empty_like(adv), torch.empty_like(adv)
    scratch = torch.empty_like(ret)
    run(g_rew * 2, adv2, scratch)
    run(torch.zeros_like(g_rew), adv0, scratch)
    lhs = (adv2 - adv).double().cpu().numpy()
    rhs = (adv - adv0).double().cpu().numpy()
    assert np.allclose(lhs, rhs, rtol=0, atol=2e-5 * np.abs(a).max())


@pytest.fixture
def rows_variant():
    """Selects a gae_rows kernel variant for one test and restores the automatic choice."""
    from tianshou_fork_b200 import _lib as L

    def set_variant(v):
        L.check(L.lib().ts_tune_set(b"gae_rows_variant", v), "ts_tune_set")

    yield set_variant
    set_variant(0)


# (E, T, buf_stride): T and strides multiples of 16 => the bulk-copy pipelined kernel is legal.
_TMA_SHAPES = [(1, 16, 16), (3, 48, 64), (5, 1024, 1024), (7, 1040, 2048), (2, 2048, 2048),
               (3, 4096, 4096), (700, 2048, 2048), (1500, 272, 288), (2, 10000, 10000)]


@pytest.mark.parametrize("variant", [2, 3, 4, 5, 6, 7, 8, 9])
@pytest.mark.parametrize("dt", ["f32", "f64"])
def test_rows_pipelined_variants_vs_oracle(variant, dt, rows_variant):
    """Every cp.async.bulk-pipelined variant on whole / partial / multi-chunk rows, more rows
    than persistent CTAs (several rows per CTA), padded buffer rows, f32 and f64 storage."""
    from tianshou_fork_b200 import _lib as L
    rows_variant(variant)
    tdt = torch.float32 if dt == "f32" else torch.float64
    for E, T, cap in _TMA_SHAPES:
        rew, v_s, v_next, term, trunc = c2_rollout(E, T, seed=E + T, p_term=1 / 50,
                                                   p_trunc=1 / 100)
        ref_ret, ref_adv = orc.episodic_return_rows(rew, v_s, v_next, term, trunc, 0.99, 0.95)
        pad = lambda a: np.concatenate([a, np.full((E, cap - T), 3, a.dtype)], axis=1)
        g_rew = torch.from_numpy(pad(rew)).cuda().to(tdt)
        g_term = torch.from_numpy(pad(term.view(np.uint8))).cuda()
        g_trunc = torch.from_numpy(pad(trunc.view(np.uint8))).cuda()
        g_vs = torch.from_numpy(v_s).cuda().to(tdt)
        g_vn = torch.from_numpy(v_next).cuda().to(tdt)
        adv = torch.full((E, T), float("nan"), dtype=tdt, device="cuda")
        ret = torch.full((E, T), float("nan"), dtype=tdt, device="cuda")
        code = L.dtype_code(g_rew)
        for _ in range(2):
            L.check(L.lib().ts_gae_rows(L.ptr(g_rew), L.ptr(g_vs), L.ptr(g_vn), L.ptr(g_term),
                                        L.ptr(g_trunc), E, T, cap, T, 0.99, 0.95, code, code,
                                        code, L.ptr(adv), L.ptr(ret), L.stream()), "ts_gae_rows")
        tol = RTOL if dt == "f32" else 1e-9
        assert _close(adv.cpu().numpy(), ref_adv, tol), (E, T, cap)
        assert _close(ret.cpu().numpy(), ref_ret, tol), (E, T, cap)
        # the padding of the buffer rows is never written through
        assert torch.all(g_rew[:, T:] == 3)


def test_rows_pipelined_matches_register_kernel_f64(rows_variant):
    """Two independent kernels (different composition order of the same affine maps) agree to
    f64 rounding noise on a C2-shaped batch."""
    from tianshou_fork_b200 import _lib as L
    E, T = 600, 2048
    rew, v_s, v_next, term, trunc = c2_rollout(E, T, seed=9)
    d = lambda a: torch.from_numpy(a).cuda()
    g = [d(rew), d(v_s), d(v_next), d(term.view(np.uint8)), d(trunc.view(np.uint8))]
    outs = []
    for variant in (1, 4):
        rows_variant(variant)
        adv = torch.empty(E, T, dtype=torch.float64, device="cuda")
        ret = torch.empty(E, T, dtype=torch.float64, device="cuda")
        L.check(L.lib().ts_gae_rows(*[L.ptr(t) for t in g], E, T, T, T, 0.99, 0.95, 0, 0, 1,
                                    L.ptr(adv), L.ptr(ret), L.stream()), "ts_gae_rows")
        outs.append((adv, ret))
    for a, b in zip(outs[0], outs[1]):
        assert float((a - b).abs().max()) <= 1e-12 * float(a.abs().max())


@pytest.mark.parametrize("mode", ["graph", "eager", "zero_copy", "overlap"])
def test_host_pipeline_matches_oracle(mode):
    """Pinned host buffers in and out (the e2e leg of bench.py): chunked copy/kernel/copy DAG
    replayed from a CUDA graph, the same DAG issued eagerly, and the zero-copy form where the
    kernel's bulk copies read and write the pinned host memory directly."""
    from tianshou_fork_b200.pipeline import GaeHostPipeline
    E, T = 1000, 512
    rew, v_s, v_next, term, trunc = c2_rollout(E, T, seed=77, p_term=1 / 60)
    ref_ret, ref_adv = orc.episodic_return_rows(rew, v_s, v_next, term, trunc, 0.99, 0.95)
    host = [torch.from_numpy(a).pin_memory() for a in
            (rew, v_s, v_next, term.view(np.uint8), trunc.view(np.uint8))]
    h_adv = torch.empty(E, T).pin_memory()
    h_ret = torch.empty(E, T).pin_memory()
    pipe = GaeHostPipeline(E, T, chunks=7)
    pipe.use_graph = mode == "graph"
    pipe.zero_copy = mode == "zero_copy"
    pipe.overlap = mode == "overlap"
    for it in range(3):  # replays reuse the captured graph; results must not depend on it
        h_adv.fill_(float("nan"))
        h_ret.fill_(float("nan"))
        pipe.run(*host, h_adv, h_ret, 0.99, 0.95)
        pipe.synchronize()
        assert _close(h_adv.numpy(), ref_adv) and _close(h_ret.numpy(), ref_ret), (mode, it)
    # new input values in the SAME buffers are picked up by the replay
    host[0].mul_(2.0)
    ref_ret2, ref_adv2 = orc.episodic_return_rows(rew * 2, v_s, v_next, term, trunc, 0.99, 0.95)
    pipe.run(*host, h_adv, h_ret, 0.99, 0.95)
    pipe.synchronize()
    assert _close(h_adv.numpy(), ref_adv2) and _close(h_ret.numpy(), ref_ret2)


def test_host_pipeline_overlapped_calls():
    """overlap=True: several passes over DIFFERENT host rollouts enqueued back to back (the
    input copies of pass k+1 run under the output copies of pass k), all correct after one
    synchronize(); the staging buffers are shared between the passes."""
    from tianshou_fork_b200.pipeline import GaeHostPipeline
    E, T, P = 300, 256, 5
    pipe = GaeHostPipeline(E, T, chunks=4)
    pipe.overlap = True
    passes = []
    for p in range(P):
        rew, v_s, v_next, term, trunc = c2_rollout(E, T, seed=900 + p, p_term=1 / 40)
        host = [torch.from_numpy(a).pin_memory() for a in
                (rew, v_s, v_next, term.view(np.uint8), trunc.view(np.uint8))]
        out = (torch.full((E, T), float("nan")).pin_memory(),
               torch.full((E, T), float("nan")).pin_memory())
        passes.append((host, out, orc.episodic_return_rows(rew, v_s, v_next, term, trunc,
                                                           0.99, 0.95)))
    for host, out, _ in passes:
        pipe.run(*host, *out, 0.99, 0.95)
    pipe.synchronize()
    for host, (h_adv, h_ret), (ref_ret, ref_adv) in passes:
        assert _close(h_adv.numpy(), ref_adv) and _close(h_ret.numpy(), ref_ret)


@pytest.mark.parametrize("E,T", [(8192, 128), (2100, 512), (1100, 1000), (5000, 250),
                                 (140000, 8), (1030, 1024), (700, 2040), (9000, 120)])
def test_rows_short_rows_packed_vs_oracle(E, T):
    """Contiguous rows shorter than a tile are walked as one sequence with a cut at every row
    end (ts_tune "gae_rows_packed"): same numbers as the oracle and as the row-per-tile path."""
    from tianshou_fork_b200 import _lib as L
    rew, v_s, v_next, term, trunc = c2_rollout(E, T, seed=E + T, p_term=1 / 300, p_trunc=1 / 900)
    ref_ret, ref_adv = orc.episodic_return_rows(rew, v_s, v_next, term, trunc, 0.99, 0.95)
    d = lambda a: torch.from_numpy(a).cuda()
    g = [d(rew), d(v_s), d(v_next), d(term.view(np.uint8)), d(trunc.view(np.uint8))]
    outs = []
    for packed in (1, 0):
        L.check(L.lib().ts_tune_set(b"gae_rows_packed", packed), "tune")
        adv = torch.full((E, T), float("nan"), device="cuda")
        ret = torch.full((E, T), float("nan"), device="cuda")
        L.check(L.lib().ts_gae_rows(*[L.ptr(t) for t in g], E, T, T, T, 0.99, 0.95, 0, 0, 0,
                                    L.ptr(adv), L.ptr(ret), L.stream()), "ts_gae_rows")
        outs.append((adv.cpu().numpy(), ret.cpu().numpy()))
    L.check(L.lib().ts_tune_set(b"gae_rows_packed", 1), "tune")
    for a, r in outs:
        assert _close(a, ref_adv) and _close(r, ref_ret)
    assert np.allclose(outs[0][0], outs[1][0], rtol=1e-6, atol=1e-6)


@pytest.mark.parametrize("case", ["ragged", "wrapped"])
def test_listing_runs_path_equals_indexed_path(case, monkeypatch):
    """sample(0) of ragged / wrapped sub-buffers: the batch is the buffer's own listing, so
    compute_episodic_return marks the unfinished tails at batch positions (ts_mark_unfinished)
    and runs the flat kernel without indices.  Same result as the indexed path (taken when the
    index tensor is not the listing itself) and as the oracle."""
    import tianshou_fork_b200 as ts
    from oracle import oracle as orc
    from tianshou_fork_b200.policy import base as policy_base
    monkeypatch.setattr(policy_base, "LISTING_MIN_N", 1 << 12)  # take the path at test size
    rng = np.random.default_rng(77)
    E, cap = 24, 4096
    buf = ts.VectorReplayBuffer(E * cap, E, rew_dtype=torch.float32)
    steps = 4096 + 700 if case == "wrapped" else 3800
    alive = np.ones(E, bool)
    for t in range(steps):
        if case == "ragged" and t % 500 == 499:  # some env streams stop contributing
            alive[rng.integers(0, E)] = False
        ids = np.flatnonzero(alive) if case == "ragged" else np.arange(E)
        m = ids.size
        buf.add(ts.Batch(obs=np.zeros((m, 1), np.float32), act=np.zeros(m, np.int64),
                         rew=rng.standard_normal(m).astype(np.float32),
                         terminated=rng.random(m) < 0.01, truncated=rng.random(m) < 0.003,
                         obs_next=np.zeros((m, 1), np.float32), info={}), buffer_ids=ids)
    batch, indices = buf.sample(0)
    n = len(batch)
    assert n >= 1 << 16 and buf.is_canonical(indices) is None
    assert buf.listing_runs(indices) is not None
    v_s = torch.randn(n, device="cuda")
    v_n = torch.randn(n, device="cuda")
    ret_a, adv_a = ts.BasePolicy.compute_episodic_return(batch, buf, indices, v_n, v_s, 0.99, 0.95)
    other = indices.clone()  # same values, not the listing tensor -> indexed kernel
    assert buf.listing_runs(other) is None
    ret_b, adv_b = ts.BasePolicy.compute_episodic_return(batch, buf, other, v_n, v_s, 0.99, 0.95)
    assert torch.equal(adv_a, adv_b) and torch.equal(ret_a, ret_b)
    idx = indices.cpu().numpy()
    want_ret, want_adv = orc.compute_episodic_return(
        batch.rew.cpu().numpy(), batch.terminated.cpu().numpy(), batch.truncated.cpu().numpy(),
        buf.terminated.cpu().numpy()[idx], idx, buf.unfinished_index(),
        v_s_=v_n.cpu().numpy(), v_s=v_s.cpu().numpy(), gamma=0.99, gae_lambda=0.95)
    assert np.allclose(adv_a.cpu().numpy(), want_adv, rtol=1e-5, atol=1e-5)
    assert np.allclose(ret_a.cpu().numpy(), want_ret, rtol=1e-5, atol=1e-5)


def test_c2_full_size_through_the_drop_in_api():
    """configs[1] through ``VectorReplayBuffer.add`` -> ``sample(0)`` ->
    ``BasePolicy.compute_episodic_return``: the canonical batch (rows kernel) and the same batch
    with a foreign index tensor (indexed flat kernel), every one of the 8 388 608 transitions
    against the oracle; NumPy in -> NumPy f64 out like the reference."""
    import tianshou_fork_b200 as ts
    E, T = 4096, 2048
    rew, v_s, v_next, term, trunc = c2_rollout(E, T, seed=1)
    buf = ts.VectorReplayBuffer(E * T, E, rew_dtype=torch.float32)
    z = np.zeros(E, np.uint8)
    for t in range(T):
        buf.add(ts.Batch(obs=z, act=z, rew=rew[:, t], terminated=term[:, t],
                         truncated=trunc[:, t], obs_next=z, info={}))
    batch, indices = buf.sample(0)
    assert buf.is_canonical(indices) == (E, T)
    ref_ret, ref_adv = orc.episodic_return_rows(rew, v_s, v_next, term, trunc, 0.99, 0.95,
                                                threads=orc.max_threads())
    g_vs, g_vn = torch.from_numpy(v_s).cuda().view(-1), torch.from_numpy(v_next).cuda().view(-1)
    ret, adv = ts.BasePolicy.compute_episodic_return(batch, buf, indices, g_vn, g_vs, 0.99, 0.95)
    assert _close(adv.cpu().numpy(), ref_adv.ravel()) and _close(ret.cpu().numpy(), ref_ret.ravel())
    other = indices.clone()
    assert buf.is_canonical(other) is None
    ret2, adv2 = ts.BasePolicy.compute_episodic_return(batch, buf, other, g_vn, g_vs, 0.99, 0.95)
    assert _close(adv2.cpu().numpy(), ref_adv.ravel())
    assert _close(ret2.cpu().numpy(), ref_ret.ravel())
    # the reference's calling convention: NumPy arrays in, float64 NumPy arrays out
    ret3, adv3 = ts.BasePolicy.compute_episodic_return(
        batch, buf, indices, v_next.ravel(), v_s.ravel(), 0.99, 0.95)
    assert isinstance(ret3, np.ndarray) and ret3.dtype == np.float64
    assert _close(adv3, ref_adv.ravel(), 1e-9) and _close(ret3, ref_ret.ravel(), 1e-9)


def test_flat_workspace_shared_by_big_and_small_calls():
    """One workspace, one stream: pipelined ranges (2^21 steps, gamma = lambda = 1, no episode
    end: every range waits for its neighbour's tagged head value), then a short sequence through
    the look-back kernel, then the big one again with other rewards.  The look-back path must not
    reset the call counter the tags derive from (a stale head value would be accepted)."""
    from tianshou_fork_b200 import _lib as L
    N, n_small = 1 << 21, 100_000
    rng = np.random.default_rng(12)
    v_s = rng.standard_normal(N).astype(np.float32)
    v_n = rng.standard_normal(N).astype(np.float32)
    z = np.zeros(N, dtype=bool)
    d = lambda a: torch.from_numpy(a).cuda()
    nbytes = L.lib().ts_gae_flat_workspace_bytes(N)
    ws = torch.zeros(nbytes, dtype=torch.uint8, device="cuda")
    adv = torch.empty(N, dtype=torch.float64, device="cuda")
    ret = torch.empty(N, dtype=torch.float64, device="cuda")

    def call(g_rew, n):
        L.check(L.lib().ts_gae_flat(L.ptr(g_rew), L.ptr(g[0]), L.ptr(g[1]), L.ptr(g[2]),
                                    L.ptr(g[2]), None, None, None, None, 0, 0, n, 1.0, 1.0, 1, 0,
                                    1, L.ptr(adv), L.ptr(ret), L.ptr(ws), nbytes, L.stream()),
                "ts_gae_flat")

    g = [d(v_s), d(v_n), d(z.view(np.uint8))]
    for rounds in range(3):
        rew = rng.standard_normal(N)
        call(d(rew), N)
        assert _close(adv.cpu().numpy(), orc.gae_return(v_s, v_n, rew, z, 1.0, 1.0), 1e-9), rounds
        rew_s = rng.standard_normal(n_small)
        call(d(rew_s), n_small)
        want = orc.gae_return(v_s[:n_small], v_n[:n_small], rew_s, z[:n_small], 1.0, 1.0)
        assert _close(adv[:n_small].cpu().numpy(), want, 1e-9), rounds


@pytest.mark.parametrize("kinds", ["pageable", "pinned", "device_batch", "f64_values"])
def test_host_call_numpy_in_numpy_out(kinds, monkeypatch):
    """The drop-in with the reference's calling convention (NumPy critic values in, float64
    NumPy out) on a canonical rollout goes through pipeline.GaeHostCall: pageable arrays, NumPy
    views of pinned memory, a batch that is already on the device, f64 values; several chunks."""
    import tianshou_fork_b200 as ts
    from tianshou_fork_b200 import pipeline
    from tianshou_fork_b200.policy import base as policy_base
    monkeypatch.setattr(policy_base, "HOST_CALL_MIN_N", 1 << 10)
    monkeypatch.setattr(pipeline.GaeHostCall, "CHUNK_BYTES", 1 << 20)  # ~10 chunks at this size
    E, T = 640, 512
    rew, v_s, v_next, term, trunc = c2_rollout(E, T, seed=21, p_term=1 / 70)
    buf = ts.VectorReplayBuffer(E * T, E, rew_dtype=torch.float32)
    z = np.zeros(E, np.uint8)
    for t in range(T):
        buf.add(ts.Batch(obs=z, act=z, rew=rew[:, t], terminated=term[:, t],
                         truncated=trunc[:, t], obs_next=z, info={}))
    batch, indices = buf.sample(0)
    ref_ret, ref_adv = orc.episodic_return_rows(rew, v_s, v_next, term, trunc, 0.99, 0.95)
    pin = lambda a: torch.from_numpy(a).pin_memory().numpy()  # noqa: E731
    vs, vn = v_s.ravel(), v_next.ravel()
    if kinds == "pinned":
        vs, vn = pin(vs), pin(vn)
        batch = ts.Batch(rew=pin(rew.ravel()), terminated=pin(term.ravel()),
                         truncated=pin(trunc.ravel()))
    elif kinds == "pageable":
        batch = ts.Batch(rew=rew.ravel().astype(np.float64), terminated=term.ravel(),
                         truncated=trunc.ravel())
    elif kinds == "f64_values":
        vs, vn = vs.astype(np.float64), vn.astype(np.float64)
    hc = pipeline.gae_host_call(buf.device)
    n0 = hc.launches
    for _ in range(2):
        ret, adv = ts.BasePolicy.compute_episodic_return(batch, buf, indices, vn, vs, 0.99, 0.95)
        assert isinstance(ret, np.ndarray) and ret.dtype == np.float64 and ret.shape == (E * T,)
        assert _close(adv, ref_adv.ravel(), 1e-9) and _close(ret, ref_ret.ravel(), 1e-9)
    assert hc.launches - n0 >= 4  # chunked
    # our extension: float32 out for NumPy callers halves the device-to-host bytes
    ret32, adv32 = ts.BasePolicy.compute_episodic_return(batch, buf, indices, vn, vs, 0.99, 0.95,
                                                         out_dtype=torch.float32)
    assert ret32.dtype == np.float32 and _close(adv32, ref_adv.ravel())
