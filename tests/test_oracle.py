"""Pins the CPU oracle (oracle/) against the reference.

Two anchors: (1) the reference's own known-answer vectors, re-typed from
test/base/test_returns.py and test/base/test_buffer.py; (2) tests/golden/*.npz, outputs of
the imported reference (tests/golden/make_golden.py).  CPU only.
"""
import numpy as np
import pytest

from oracle import oracle as orc
from conftest import load_golden


# ---- reference known answers: test/base/test_returns.py:22-112 ---------------------------
def _episodic(done, trunc, rew, v=None, gamma=0.1, lam=1.0, unfinished_last=True):
    done = np.array(done, dtype=bool)
    trunc = np.array(trunc, dtype=bool) if trunc is not None else np.zeros_like(done)
    term = done & ~trunc
    rew = np.array(rew, dtype=np.float64)
    n = len(rew)
    idx = np.arange(n)
    end = term | trunc
    unfinished = np.array([n - 1] if not end[-1] else [], dtype=np.int64)
    v_s_ = None if v is None else np.array(v, dtype=np.float64)
    return orc.compute_episodic_return(rew, term, trunc, term, idx, unfinished, v_s_=v_s_,
                                       gamma=gamma, gae_lambda=lam)


def test_known_answer_mc_returns():
    # test_returns.py:25-38
    ret, _ = _episodic([1, 0, 0, 1, 0, 1, 0, 1.0], None, [0, 1, 2, 3, 4, 5, 6, 7.0],
                       v=np.zeros(8))
    assert np.allclose(ret, [0, 1.23, 2.3, 3, 4.5, 5, 6.7, 7])
    # :39-51
    ret, _ = _episodic([0, 1, 0, 1, 0, 1, 0.0], None, [7, 6, 1, 2, 3, 4, 5.0], v=np.zeros(7))
    assert np.allclose(ret, [7.6, 6, 1.2, 2, 3.4, 4, 5])
    # :52-65 (last episode unfinished)
    ret, _ = _episodic([0, 1, 0, 1, 0, 0, 1.0], None, [7, 6, 1, 2, 3, 4, 5.0], v=np.zeros(7))
    assert np.allclose(ret, [7.6, 6, 1.2, 2, 3.45, 4.5, 5])


def test_known_answer_gae():
    # test_returns.py:67-83: gamma .99 lambda .95 with terminated episodes
    done = [0, 0, 0, 1.0, 0, 0, 0, 1, 0, 0, 0, 1]
    rew = [101, 102, 103.0, 200, 104, 105, 106, 201, 107, 108, 109, 202]
    v = np.array([2.0, 3.0, 4, -1, 5.0, 6.0, 7, -2, 8.0, 9.0, 10, -3])
    ret, _ = _episodic(done, None, rew, v=v, gamma=0.99, lam=0.95)
    expect = [454.8344, 376.1143, 291.298, 200.0, 464.5610, 383.1085, 295.387, 201.0,
              474.2876, 390.1027, 299.476, 202.0]
    assert np.allclose(ret, expect)
    # :85-112: same but the three ends are truncations -> bootstrap through v
    trunc_b = np.array([0, 0, 0, 1, 0, 0, 0, 1, 0, 0, 0, 0], dtype=bool)
    term_b = np.array([0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 1], dtype=bool)
    ret, _ = orc.compute_episodic_return(
        np.array(rew), term_b, trunc_b, term_b, np.arange(12),
        np.array([], dtype=np.int64), v_s_=v, gamma=0.99, gae_lambda=0.95)
    expect = [454.0109, 375.2386, 290.3669, 199.01, 462.9138, 381.3571, 293.5248, 199.02,
              474.2876, 390.1027, 299.476, 202.0]
    assert np.allclose(ret, expect)


# ---- reference known answers: test/base/test_buffer.py:497-594 ---------------------------
def test_known_answer_segtree_corner():
    t = orc.SegTree(8)
    t[np.array([0, 2])] = np.array([0.1, 0.1])  # naive = [.1, 0, .1, 0...]
    assert np.allclose(t.reduce(), 0.2)
    idx = t.get_prefix_sum_idx(np.array([0, 0.1, 0.1 + 1e-6, 0.2 - 1e-6]))
    assert np.all(idx == [0, 0, 2, 2])  # test_buffer.py:561-565
    with pytest.raises(AssertionError):
        t.get_prefix_sum_idx(0.2)  # :566-567
    with pytest.raises(IndexError):
        t[8]  # noqa  (:504-505: beyond the heap array)


def test_segtree_vs_naive():
    rng = np.random.default_rng(0)
    for size in (8, 100, 16384):
        t = orc.SegTree(size)
        naive = np.zeros(size)
        for _ in range(50):
            idx = rng.integers(0, size, 64)
            val = rng.random(64)
            t[idx] = val
            for i, v in zip(idx, val):  # sequential: last wins
                naive[i] = v
        assert np.allclose(t.reduce(), naive.sum())
        for _ in range(20):
            a, b = sorted(rng.integers(0, size + 1, 2))
            if a < b:
                assert np.allclose(t.reduce(int(a), int(b)), naive[a:b].sum())
        s = rng.random(256) * t.reduce()
        idx = t.get_prefix_sum_idx(s.copy())
        cs = np.cumsum(naive)
        assert np.all(cs[idx] >= s - 1e-9)
        assert np.all(np.where(idx > 0, cs[idx - 1], 0.0) <= s + 1e-9)


# ---- golden fixtures -----------------------------------------------------------------------
def test_golden_gae():
    g = load_golden("gae")
    assert set(g) >= {"small", "wrap", "ragged", "long4096", "mc4096", "noend"}
    for name, c in g.items():
        idx = c["indices"]
        ret, adv = orc.compute_episodic_return(
            c["buf_rew"][idx], c["buf_terminated"][idx], c["buf_truncated"][idx],
            c["buf_terminated"][idx], idx, c["unfinished"], v_s_=c["v_s_"], v_s=c["v_s"],
            gamma=float(c["gamma"]), gae_lambda=float(c["lam"]))
        # same f64 operation order => bit-identical
        assert np.array_equal(adv, c["adv"]), name
        assert np.array_equal(ret, c["returns"]), name
    c = g["small"]
    idx = c["indices"]
    args = (c["buf_rew"][idx], c["buf_terminated"][idx], c["buf_truncated"][idx],
            c["buf_terminated"][idx], idx, c["unfinished"])
    ret, adv = orc.compute_episodic_return(*args, v_s_=c["v_s_"], v_s=None,
                                           gamma=float(c["gamma"]), gae_lambda=float(c["lam"]))
    assert np.array_equal(ret, c["returns_roll"]) and np.array_equal(adv, c["adv_roll"])
    ret, adv = orc.compute_episodic_return(*args, v_s_=None, v_s=None,
                                           gamma=float(c["gamma"]), gae_lambda=1.0)
    assert np.array_equal(ret, c["returns_mc"]) and np.array_equal(adv, c["adv_mc"])


def test_rows_form_matches_flat_form():
    """orc_episodic_return_rows == compute_episodic_return on an un-wrapped [E,T] buffer."""
    c = load_golden("gae")["long4096"]
    E, T = int(c["n_env"]), int(c["cap"])
    idx = c["indices"]
    assert np.array_equal(idx, np.arange(E * T))
    ret, adv = orc.episodic_return_rows(
        c["buf_rew"].reshape(E, T).astype(np.float32), c["v_s"].reshape(E, T),
        c["v_s_"].reshape(E, T), c["buf_terminated"].reshape(E, T),
        c["buf_truncated"].reshape(E, T), float(c["gamma"]), float(c["lam"]), threads=2)
    # rew is rounded to f32 in the rows form: compare against the flat oracle on the same
    ret2, adv2 = orc.compute_episodic_return(
        c["buf_rew"].astype(np.float32).astype(np.float64), c["buf_terminated"],
        c["buf_truncated"], c["buf_terminated"], idx, c["unfinished"], v_s_=c["v_s_"],
        v_s=c["v_s"], gamma=float(c["gamma"]), gae_lambda=float(c["lam"]))
    assert np.array_equal(adv.ravel(), adv2) and np.array_equal(ret.ravel(), ret2)


def test_golden_index():
    g = load_golden("index")
    for name, c in g.items():
        if name.startswith("s_"):
            size, last = int(c["size"]), int(c["last_index"][0])
            assert np.array_equal(orc.prev_single(c["query"], c["done"], last, size), c["prev"])
            assert np.array_equal(orc.next_single(c["query"], c["done"], last, size), c["next"])
        else:
            args = (c["offset"], c["done"], c["last_index"], c["lengths"])
            assert np.array_equal(orc.prev_index(c["query"], *args), c["prev"]), name
            assert np.array_equal(orc.next_index(c["query"], *args), c["next"]), name


def test_ringstate_matches_reference_bookkeeping():
    """RingState.add replays the same add sequence as make_golden.fill_vecbuf."""
    g = load_golden("index")
    for i, (name, n_env, total, steps, ragged) in enumerate(
            [("v_partial", 4, 20, 11, True), ("v_wrap", 4, 20, 57, False),
             ("v_big", 16, 1024, 700, True), ("v_one", 1, 16, 40, False)]):
        rng = np.random.default_rng(200 + i)
        st = orc.RingState(total, n_env)
        for _ in range(steps):
            ids = np.arange(n_env)
            if ragged:
                ids = ids[rng.random(n_env) < 0.7]
                if ids.size == 0:
                    continue
            m = ids.size
            term = rng.random(m) < 0.15
            trunc = (rng.random(m) < 0.05) & ~term
            rng.standard_normal((m, 3)); rng.integers(0, 4, m)
            rew = rng.standard_normal(m)
            rng.standard_normal((m, 3))
            st.add(rew, term, trunc, ids)
        c = g[name]
        assert np.array_equal(st.done, c["done"]), name
        assert np.array_equal(st.last_index, c["last_index"])
        assert np.array_equal(st.lengths, c["lengths"])
        assert np.array_equal(st.index, c["ring_index"])
        assert np.array_equal(st.unfinished_index(), c["unfinished"])
        assert np.array_equal(st.sample_indices_all(), c["all_indices"])


def test_golden_nstep():
    g = load_golden("nstep")
    for name, c in g.items():
        n_env = int(c["n_env"])
        st = orc.RingState(c["done"].size, n_env)
        st.done, st.terminated, st.truncated = c["done"], c["terminated"], c["truncated"]
        st.rew, st.last_index, st.lengths = c["rew"], c["last_index"], c["lengths"]
        st.index = c["ring_index"]
        st.size = c["lengths"]
        table = c["table"]
        D = table.shape[1]
        fn = (lambda t: table[t]) if D > 1 else (lambda t: table[t].ravel())
        out = orc.compute_nstep_return(st, c["indice"], fn, float(c["gamma"]), int(c["n_step"]))
        # reference casts the f64 result to the f32 target dtype (policy/base.py:437)
        assert np.array_equal(out.astype(np.float32), c["returns"]), name


def test_golden_segtree():
    g = load_golden("segtree")
    for name, c in g.items():
        t = orc.SegTree(int(c["size"]))
        for idx, val in zip(c["idx_log"], c["val_log"]):
            t[idx] = val
        assert np.array_equal(t._value, c["tree"]), name
        s = c["scalar"].copy()
        assert np.array_equal(t.get_prefix_sum_idx(s), c["prefix_idx"])
        assert np.array_equal(s, c["scalar_after"])  # in-place mutation (segtree.py:134)
        s32 = c["scalar32"].copy()
        assert np.array_equal(t.get_prefix_sum_idx(s32), c["prefix_idx32"])
        assert np.array_equal(s32, c["scalar32_after"])
        red = [t.reduce(int(a), int(b)) for a, b in c["reduce_se"]]
        assert np.array_equal(np.array(red), c["reduce"])
        assert t.reduce(1, -1) == float(c["reduce_neg_end"])


def test_golden_per():
    g = load_golden("per")
    for name in ("pv", "pv_nonorm"):
        c = g[name]
        # Reference quirk: the vector PER manager re-runs PrioritizedReplayBuffer.__init__
        # with buffer_list[0].options (manager.py:211-214), which never carries
        # weight_norm (prio.py:39), so the manager ALWAYS normalises, even though
        # "pv_nonorm" was constructed with weight_norm=False.
        per = orc.PER(int(c["size"]), 0.6, 0.4, weight_norm=True)
        per.weight._value[:] = c["tree_after_add"]
        for it in range(c["u"].shape[0]):
            idx = per.sample_indices(c["u"][it])
            w = per.batch_weight(idx)
            gold_idx = c["idx"][it].copy()
            if it % 2 == 0:
                gold_idx[-4:] = idx[-4:]  # generator overwrote the tail after sampling
            assert np.array_equal(idx, gold_idx), (name, it)
            assert np.allclose(w, c["weight"][it], rtol=1e-12)
            per.update_weight(c["idx"][it], c["td"][it])
            assert np.array_equal(per.weight._value, c["tree"][it]), (name, it)
            assert per._max_prio == c["max_prio"][it] and per._min_prio == c["min_prio"][it]


def test_golden_ppo_loss():
    g = load_golden("ppo")
    x = {k: v.astype(np.float64) for k, v in g["in"].items()}
    for name, c in g.items():
        if name == "in":
            continue
        norm_adv, value_clip, dual, eps_clip, _ = c["cfg"]
        cl, vl, dlogp, dvalue, adv = orc.ppo_loss(
            x["logp"], x["logp_old"], x["adv"], x["returns"], x["v_s"], x["value"],
            eps_clip=eps_clip, dual_clip=(dual or None), value_clip=bool(value_clip),
            norm_adv=bool(norm_adv))
        assert np.allclose(cl, c["clip_loss"], rtol=1e-12), name
        assert np.allclose(vl, c["vf_loss"], rtol=1e-12), name
        assert np.allclose(adv, c["adv_used"], rtol=1e-12, atol=1e-15)
        if not norm_adv:  # with norm_adv autograd also flows through mean/std: adv is an input
            assert np.allclose(dlogp, c["dlogp"], rtol=1e-10, atol=1e-15), name
        assert np.allclose(dvalue, c["dvalue"] / 0.5, rtol=1e-10, atol=1e-15), name


def test_golden_stack_gather():
    """Frame-stack gathers vs the reference's buf[q] (make_golden.gen_stack; base.py:317-378)."""
    g = load_golden("stack")
    for name, c in g.items():
        if name == "ring":
            size = int(c["size"])
            args = (np.array([0, size]), c["done"], c["last_index"], np.array([size]))
            assert np.array_equal(orc.stack_gather(c["frames"], c["query"], 3, *args), c["obs"])
            assert np.array_equal(orc.stack_gather(c["frames_next"], c["query"], 3, *args),
                                  c["obs_next"])
            continue
        args = (c["offset"], c["done"], c["last_index"], c["lengths"])
        k = c["obs"].shape[1]
        assert np.array_equal(orc.stack_gather(c["frames"], c["query"], k, *args), c["obs"]), name
        assert np.array_equal(orc.stack_gather(c["frames"], c["query"], k, *args, from_next=True),
                              c["obs_next"]), name
        assert np.array_equal(orc.stack_gather(c["frames"], c["query"], 2, *args), c["get2"]), name


# ---- BASELINE.json configs[3] at full size: oracle vs the reference's own outputs -----------------
def test_oracle_c4_full_size_vs_reference_golden():
    """tests/golden/c4.npz (make_golden.gen_c4: the reference's PrioritizedVectorReplayBuffer of
    2^20 slots in 16 wrapped rings): prev / next, the PER sampled indices, importance weights and
    the n-step targets (n = 3, D = 1 and 51) of batches 32 / 256 / 4096 from the oracle."""
    import hashlib
    import helpers as H
    g = load_golden("c4")["c4"]
    rew, term, trunc = H.c4_stream()
    s_rew, s_term, s_trunc, ring_index, last = H.c4_final_storage(rew, term, trunc)
    sha = lambda a: np.frombuffer(hashlib.sha256(np.ascontiguousarray(a).tobytes()).digest(),  # noqa: E731
                                  np.uint8)
    assert np.array_equal(sha(s_rew), g["rew_sha"])
    assert np.array_equal(sha(s_term | s_trunc), g["done_sha"])
    E, sub = H.C4_E, H.C4_SUB
    st = orc.RingState(E * sub, E)
    st.rew, st.terminated, st.truncated, st.done = s_rew, s_term, s_trunc, s_term | s_trunc
    st.index[:], st.size[:], st.lengths[:] = ring_index, sub, sub
    st.last_index = last
    assert np.array_equal(last, g["last_index"]) and np.array_equal(st.lengths, g["lengths"])
    assert np.array_equal(st.unfinished_index(), g["unfinished"])
    assert np.array_equal(st.prev(g["q"]), g["prev"]) and np.array_equal(st.next(g["q"]), g["next"])
    per = orc.PER(E * sub, 0.6, 0.4)
    per.weight[np.arange(E * sub)] = H.c4_leaves()
    assert per.weight.reduce() == float(g["root0"])
    rng = np.random.default_rng(4002)
    rng.integers(0, E * sub, 4096)  # the generator drew q first
    for bsz in H.C4_BATCHES:
        np.random.seed(2000 + bsz)
        idx = per.sample_indices(np.random.rand(bsz))
        assert np.array_equal(idx, g[f"b{bsz}/idx"]), bsz  # bit-exact
        assert np.allclose(per.batch_weight(idx), g[f"b{bsz}/weight"], rtol=1e-12)
        for D in H.C4_DIMS:
            r = orc.compute_nstep_return(st, idx, lambda t: H.c4_target_q(t, D), 0.99, 3)
            r = r.astype(np.float32) if D > 1 else r.astype(np.float32).reshape(bsz, 1)
            assert np.array_equal(r[:64].reshape(g[f"b{bsz}/d{D}/returns_head"].shape),
                                  g[f"b{bsz}/d{D}/returns_head"])
            assert np.array_equal(sha(r), g[f"b{bsz}/d{D}/returns_sha"]), (bsz, D)
        td = rng.standard_normal(bsz).astype(np.float32)
        idx_u = idx.copy()
        k = bsz // 20
        idx_u[:k] = idx_u[k:2 * k]
        per.update_weight(idx_u, td)
        uniq = np.unique(idx_u)
        # NumPy computed these very leaves (same library): bit-equal here, 6e-7 on the GPU
        assert np.array_equal(per.weight[uniq], g[f"b{bsz}/leaves"])
        assert per.weight.reduce() == float(g[f"b{bsz}/root"])
        assert np.array_equal([per._max_prio, per._min_prio], g[f"b{bsz}/prio"])


def test_numba_port_matches_c_oracle_and_goldens():
    """oracle/oracle_nb.py (the timed CPU baseline of bench.py: the reference's hot path restated
    in numba + NumPy) returns what the C oracle and the reference's goldens hold."""
    from oracle import oracle_nb as nb
    import helpers as H
    nb.warm()
    rng = np.random.default_rng(5)
    # GAE over a flattened rotated batch with unfinished tails
    B, cap = 7, 50
    idx = np.concatenate([k * cap + (np.arange(cap) + 3 * k) % cap for k in range(B)])
    n = idx.size
    buf_term = rng.random(B * cap) < 0.05
    term, trunc = buf_term[idx], (rng.random(n) < 0.02) & ~buf_term[idx]
    unfinished = np.array([k * cap + (3 * k - 1) % cap for k in range(0, B, 2)])
    rew, v_s, v_n = rng.standard_normal(n), rng.standard_normal(n), rng.standard_normal(n)
    want = orc.compute_episodic_return(rew, term, trunc, buf_term[idx], idx, unfinished,
                                       v_s_=v_n, v_s=v_s, gamma=0.99, gae_lambda=0.95)
    got = nb.compute_episodic_return(rew, term, trunc, buf_term, idx, unfinished, v_n, v_s,
                                     0.99, 0.95)
    assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
    # PER + n-step on the C4 golden, batch 256
    g = load_golden("c4")["c4"]
    r, t, tr = H.c4_stream()
    s_rew, s_term, s_trunc, ring_index, last = H.c4_final_storage(r, t, tr)
    size = H.C4_E * H.C4_SUB
    per = nb.PER(size, 0.6, 0.4)
    per.set_leaves(np.arange(size), H.c4_leaves())
    for bsz in (32, 256):
        np.random.seed(2000 + bsz)
        idx = per.sample_indices(bsz)
        assert np.array_equal(idx, g[f"b{bsz}/idx"])
        assert np.allclose(per.batch_weight(idx), g[f"b{bsz}/weight"], rtol=1e-12)
        off = np.arange(H.C4_E + 1) * H.C4_SUB
        lens = np.full(H.C4_E, H.C4_SUB)
        ret = nb.compute_nstep_return(
            s_rew, s_term | s_trunc, s_term, off, last, lens, g["unfinished"], idx,
            lambda tt: H.c4_target_q(tt, 51), 0.99, 3).astype(np.float32)
        assert np.array_equal(ret[:64], g[f"b{bsz}/d51/returns_head"])
        if bsz == 32:  # keep the tree in the golden's state for the next batch
            rng2 = np.random.default_rng(4002)
            rng2.integers(0, size, 4096)
            td = rng2.standard_normal(bsz).astype(np.float32)
            idx_u = idx.copy()
            idx_u[:1] = idx_u[1:2]
            per.update_weight(idx_u, td)
            assert per._value[1] == float(g["b32/root"])
