"""Sharded PER on one GPU: the G ranks are emulated as G local trees on cuda:0 (the kernels of
different ranks never wait on each other, so running them back to back is exact).  The union of
the ranks' samples must equal the reference's samples from ONE global tree, bit for bit."""
import numpy as np
import pytest
import torch

from oracle import oracle as orc

pytestmark = pytest.mark.gpu


def _fill(tree_t, stamp, counter, err, leaves, L):
    n = leaves.size
    idx = torch.arange(n, device="cuda")
    val = torch.from_numpy(leaves).cuda()
    L.check(L.lib().ts_segtree_update(L.ptr(tree_t), n, n, L.ptr(idx), L.ptr(val), L.TS_F64, 1, n,
                                      L.ptr(stamp), L.ptr(counter), L.ptr(err), L.stream()),
            "fill")


@pytest.mark.parametrize("G,cap,bsz", [(2, 1 << 13, 512), (4, 1 << 14, 4096), (8, 1 << 16, 4096),
                                       (8, 1 << 20, 4096), (1, 1 << 14, 300)])
def test_emulated_ranks_match_global_tree(G, cap, bsz):
    from tianshou_fork_b200 import _lib as L
    if G * cap > (1 << 21):
        rng = np.random.default_rng(G)
        steps = 2
    else:
        rng = np.random.default_rng(G + cap)
        steps = 4
    total = G * cap
    leaves = (np.abs(rng.standard_normal(total)) + 1e-7) ** 0.6
    glob = orc.PER(total, 0.6, 0.4)
    glob.weight[np.arange(total)] = leaves
    ranks = []
    for r in range(G):
        d = dict(tree=torch.zeros(2 * cap, dtype=torch.float64, device="cuda"),
                 stamp=torch.zeros(cap, dtype=torch.int32, device="cuda"),
                 counter=torch.zeros(4, dtype=torch.int32, device="cuda"),
                 err=torch.zeros(1, dtype=torch.int32, device="cuda"),
                 prio=torch.ones(2, dtype=torch.float64, device="cuda"),
                 wmax=torch.zeros(1, dtype=torch.float64, device="cuda"))
        _fill(d["tree"], d["stamp"], d["counter"], d["err"], leaves[r * cap:(r + 1) * cap], L)
        ranks.append(d)
    lib = L.lib()
    for step in range(steps):
        # ---- the exchange: {root, max_prio, min_prio} of every rank --------------------------
        roots3 = torch.stack([torch.cat([d["tree"][1:2], d["prio"]]) for d in ranks]).contiguous()
        gmax, gmin = roots3[:, 1].max(), roots3[:, 2].min()
        for d in ranks:
            d["prio"][0], d["prio"][1] = gmax, gmin
        u = rng.random(bsz)
        want = glob.sample_indices(u)
        want_w = glob.batch_weight(want)  # (p/min)^-beta / max over the batch (prio.py:76,101)
        ud = torch.from_numpy(u).cuda()
        idxs, ws = [], []
        for r, d in enumerate(ranks):
            idx = torch.empty(bsz, dtype=torch.int64, device="cuda")
            w = torch.empty(bsz, dtype=torch.float64, device="cuda")
            L.check(lib.ts_per_sample_sharded(
                L.ptr(d["tree"]), cap, L.ptr(ud), bsz, L.ptr(roots3), G, r, L.ptr(d["prio"]), 0.4,
                1, L.ptr(idx), L.ptr(w), L.ptr(d["wmax"]), L.ptr(d["err"]), L.stream()), "sample")
            idxs.append(idx)
            ws.append(w)
        own = torch.stack([i >= 0 for i in idxs])
        assert bool((own.sum(0) == 1).all())  # every query owned by exactly one rank
        got = sum(torch.where(i >= 0, i + r * cap, torch.zeros_like(i))
                  for r, i in enumerate(idxs)).cpu().numpy()
        assert np.array_equal(got, want), (G, step)  # bit-exact global indices
        # MAX all-reduce of the per-rank maxima, then the normalisation kernel
        wmax = torch.stack([d["wmax"] for d in ranks]).max().reshape(1)
        for w in ws:
            L.check(lib.ts_per_normalize(L.ptr(w), bsz, L.ptr(wmax), L.stream()), "norm")
        got_w = sum(ws).cpu().numpy()  # foreign entries are 0
        assert np.allclose(got_w, want_w, rtol=1e-12)
        # ---- update: every rank gets the full batch with -1 for foreign samples ---------------
        td = rng.standard_normal(bsz).astype(np.float32)
        gidx = want.copy()
        gidx[: bsz // 8] = gidx[bsz // 8: 2 * (bsz // 8)]  # duplicated indices
        glob.update_weight(gidx, td)
        tdd = torch.from_numpy(td).cuda()
        for r, d in enumerate(ranks):
            loc = np.where((gidx >= r * cap) & (gidx < (r + 1) * cap), gidx - r * cap, -1)
            locd = torch.from_numpy(loc).cuda()
            L.check(lib.ts_per_update(L.ptr(d["tree"]), cap, cap, L.ptr(locd), L.ptr(tdd), L.TS_F32,
                                      bsz, 0.6, L.ptr(d["prio"]), L.ptr(d["stamp"]),
                                      L.ptr(d["counter"]), L.ptr(d["err"]), L.stream()), "update")
            assert int(d["err"].item()) == 0  # -1 is a sentinel, not a range error
        # priority extrema: max / min over the ranks == the reference's single scalars
        assert max(float(d["prio"][0]) for d in ranks) == glob._max_prio
        assert min(float(d["prio"][1]) for d in ranks) == glob._min_prio
        # leaves through pow: tolerance, then re-anchor on the oracle's leaves (as in the
        # unsharded test) so the next round tests bit-exactness again
        for r, d in enumerate(ranks):
            got_l = d["tree"][cap:].cpu().numpy()
            ref_l = glob.weight._value[total + r * cap: total + (r + 1) * cap]
            assert np.allclose(got_l, ref_l, rtol=6e-7, atol=0)
            ch = np.nonzero(got_l != ref_l)[0]
            if ch.size:
                ci = torch.from_numpy(ch).cuda()
                cv = torch.from_numpy(ref_l[ch]).cuda()
                L.check(lib.ts_segtree_update(L.ptr(d["tree"]), cap, cap, L.ptr(ci), L.ptr(cv),
                                              L.TS_F64, 1, ch.size, L.ptr(d["stamp"]),
                                              L.ptr(d["counter"]), L.ptr(d["err"]), L.stream()),
                        "anchor")
            # shard subtree == the global tree's subtree under node G + r, every level
            node, width = G + r, 1
            loc = d["tree"].cpu().numpy()
            lnode = 1
            while lnode < 2 * cap:
                assert np.array_equal(loc[lnode:lnode + width],
                                      glob.weight._value[node:node + width])
                node, lnode, width = node * 2, lnode * 2, width * 2


def test_world_size_one_class_equals_unsharded():
    """No process group: the sharded class degenerates to PrioritizedVectorReplayBuffer."""
    import tianshou_fork_b200 as ts
    a = ts.ShardedPrioritizedVectorReplayBuffer(1 << 12, 8, alpha=0.6, beta=0.4)
    b = ts.PrioritizedVectorReplayBuffer(1 << 12, 8, alpha=0.6, beta=0.4)
    rng = np.random.default_rng(0)
    for t in range(40):
        batch = dict(obs=rng.standard_normal((8, 3)).astype(np.float32), act=rng.integers(0, 4, 8),
                     rew=rng.standard_normal(8), terminated=rng.random(8) < 0.1,
                     truncated=np.zeros(8, bool),
                     obs_next=rng.standard_normal((8, 3)).astype(np.float32), info={})
        a.add(ts.Batch(**batch))
        b.add(ts.Batch(**batch))
    for _ in range(3):
        u = torch.from_numpy(rng.random(64)).cuda()
        ia, ib = a.sample_with_uniforms(u), b.sample_with_uniforms(u)
        assert torch.equal(ia, ib) and torch.equal(a._cached_weight[1], b._cached_weight[1])
        td = torch.from_numpy(rng.standard_normal(64).astype(np.float32)).cuda()
        a.update_weight(ia, td)
        b.update_weight(ib, td)
        assert torch.equal(a.weight._value, b.weight._value)
        assert a._max_prio == b._max_prio and a._min_prio == b._min_prio
    assert torch.equal(a.to_global(ia), ia)
    # The p2p exchange keeps its own sequence words (`_p2p_state`); it must never shadow
    # ReplayBuffer._state(), which prev / next / n-step / GAE read (round-1 advisor finding).
    a._p2p_state = torch.zeros(2, dtype=torch.int64, device="cuda")  # what _setup_p2p creates
    assert callable(a._state) and callable(type(a)._state)
    q = torch.from_numpy(rng.integers(0, 1 << 12, 64)).cuda()
    assert torch.equal(a.next(q), b.next(q)) and torch.equal(a.prev(q), b.prev(q))
    tq = lambda buf, idx: torch.ones(idx.numel(), device="cuda")  # noqa: E731
    ra = ts.BasePolicy.compute_nstep_return(ts.Batch(), a, ia, tq, 0.9, 3).returns
    rb = ts.BasePolicy.compute_nstep_return(ts.Batch(), b, ib, tq, 0.9, 3).returns
    assert torch.equal(ra, rb)


@pytest.mark.parametrize("G,cap,bsz", [(2, 1 << 13, 700), (8, 1 << 15, 4096)])
def test_p2p_protocol_emulated_on_one_gpu(G, cap, bsz):
    """The in-kernel exchange (ts_per_publish -> slot buffers -> ts_per_sample_p2p) with the G
    ranks' slot buffers all on cuda:0.  Ranks run back to back, which is exact as long as no
    kernel has to wait for a LATER one: the records are published before anyone samples, and the
    weight-maximum exchange (which is mutual) is off here -- it is covered on real GPUs by
    scripts/sharded_check.py."""
    from tianshou_fork_b200 import _lib as L
    lib = L.lib()
    rng = np.random.default_rng(G * 31 + 1)
    total = G * cap
    leaves = (np.abs(rng.standard_normal(total)) + 1e-7) ** 0.6
    glob = orc.PER(total, 0.6, 0.4)
    glob.weight[np.arange(total)] = leaves
    ranks = []
    for r in range(G):
        d = dict(tree=torch.zeros(2 * cap, dtype=torch.float64, device="cuda"),
                 stamp=torch.zeros(cap, dtype=torch.int32, device="cuda"),
                 counter=torch.zeros(4, dtype=torch.int32, device="cuda"),
                 err=torch.zeros(1, dtype=torch.int32, device="cuda"),
                 prio=torch.ones(2, dtype=torch.float64, device="cuda"),
                 slots=torch.zeros(16 * G, dtype=torch.float64, device="cuda"),
                 state=torch.zeros(2, dtype=torch.int64, device="cuda"))
        _fill(d["tree"], d["stamp"], d["counter"], d["err"], leaves[r * cap:(r + 1) * cap], L)
        ranks.append(d)
    peers = torch.tensor([d["slots"].data_ptr() for d in ranks], dtype=torch.int64, device="cuda")

    def publish_all():
        for r, d in enumerate(ranks):
            L.check(lib.ts_per_publish(L.ptr(d["tree"]), L.ptr(d["prio"]), L.ptr(d["state"]),
                                       L.ptr(peers), G, r, L.stream()), "publish")

    publish_all()
    for step in range(4):
        u = rng.random(bsz)
        want = glob.sample_indices(u)
        want_w = glob.get_weight(want)  # raw (p / min_prio) ** -beta, prio.py:76
        ud = torch.from_numpy(u).cuda()
        idxs, ws = [], []
        for r, d in enumerate(ranks):
            idx = torch.empty(bsz, dtype=torch.int64, device="cuda")
            w = torch.empty(bsz, dtype=torch.float64, device="cuda")
            L.check(lib.ts_per_sample_p2p(
                L.ptr(d["tree"]), cap, L.ptr(ud), bsz, L.ptr(d["slots"]), L.ptr(peers),
                L.ptr(d["state"]), G, r, L.ptr(d["prio"]), 0.4, 0, L.ptr(idx), L.ptr(w),
                L.ptr(d["err"]), L.stream()), "sample_p2p")
            idxs.append(idx)
            ws.append(w)
            assert int(d["err"].item()) == 0  # no peer timeout, no range error
            assert d["state"].tolist() == [step + 1, step + 1]
            # the global extrema were adopted into this rank's scalars
            assert float(d["prio"][0]) == glob._max_prio and float(d["prio"][1]) == glob._min_prio
        got = sum(torch.where(i >= 0, i + r * cap, torch.zeros_like(i))
                  for r, i in enumerate(idxs)).cpu().numpy()
        assert np.array_equal(got, want), (G, step)
        assert np.allclose(sum(ws).cpu().numpy(), want_w, rtol=1e-12)
        td = rng.standard_normal(bsz).astype(np.float32)
        glob.update_weight(want, td)
        tdd = torch.from_numpy(td).cuda()
        for r, d in enumerate(ranks):
            L.check(lib.ts_per_update(L.ptr(d["tree"]), cap, cap, L.ptr(idxs[r]), L.ptr(tdd),
                                      L.TS_F32, bsz, 0.6, L.ptr(d["prio"]), L.ptr(d["stamp"]),
                                      L.ptr(d["counter"]), L.ptr(d["err"]), L.stream()), "update")
            # re-anchor the touched leaves on the oracle's f32 powers (see the unsharded test)
            loc = idxs[r][idxs[r] >= 0].unique()
            if loc.numel():
                vals = torch.from_numpy(glob.weight[loc.cpu().numpy() + r * cap]).cuda()
                L.check(lib.ts_segtree_update(L.ptr(d["tree"]), cap, cap, L.ptr(loc), L.ptr(vals),
                                              L.TS_F64, 1, loc.numel(), L.ptr(d["stamp"]),
                                              L.ptr(d["counter"]), L.ptr(d["err"]), L.stream()),
                        "anchor")
        publish_all()


@pytest.mark.parametrize("G,cap,bsz", [(8, 1 << 15, 3 * 4096 + 37), (2, 1 << 13, 28800)])
def test_p2p_owned_list_emulated_on_one_gpu(G, cap, bsz):
    """Batches larger than one update pass: ts_per_sample_p2p_owned lists the batch positions a
    rank owns (second case: more per CTA than its shared-memory list holds), and
    ts_per_update_publish_owned, scanning only those, leaves exactly the tree that the plain
    ts_per_update leaves after scanning the whole batch."""
    from tianshou_fork_b200 import _lib as L
    lib = L.lib()
    rng = np.random.default_rng(G * 17 + 5)
    total = G * cap
    leaves = (np.abs(rng.standard_normal(total)) + 1e-7) ** 0.6
    glob = orc.PER(total, 0.6, 0.4)
    glob.weight[np.arange(total)] = leaves
    ranks = []
    for r in range(G):
        d = dict(tree=torch.zeros(2 * cap, dtype=torch.float64, device="cuda"),
                 stamp=torch.zeros(cap, dtype=torch.int32, device="cuda"),
                 counter=torch.zeros(4, dtype=torch.int32, device="cuda"),
                 err=torch.zeros(1, dtype=torch.int32, device="cuda"),
                 prio=torch.ones(2, dtype=torch.float64, device="cuda"),
                 slots=torch.zeros(16 * G, dtype=torch.float64, device="cuda"),
                 state=torch.zeros(2, dtype=torch.int64, device="cuda"),
                 owned=torch.zeros(4 + bsz, dtype=torch.int32, device="cuda"))
        _fill(d["tree"], d["stamp"], d["counter"], d["err"], leaves[r * cap:(r + 1) * cap], L)
        ranks.append(d)
    peers = torch.tensor([d["slots"].data_ptr() for d in ranks], dtype=torch.int64, device="cuda")
    for r, d in enumerate(ranks):
        L.check(lib.ts_per_publish(L.ptr(d["tree"]), L.ptr(d["prio"]), L.ptr(d["state"]),
                                   L.ptr(peers), G, r, L.stream()), "publish")
    for step in range(3):
        u = rng.random(bsz)
        want = glob.sample_indices(u)
        ud = torch.from_numpy(u).cuda()
        td = rng.standard_normal(bsz).astype(np.float32)
        tdd = torch.from_numpy(td).cuda()
        idxs = []
        for r, d in enumerate(ranks):
            idx = torch.empty(bsz, dtype=torch.int64, device="cuda")
            w = torch.empty(bsz, dtype=torch.float64, device="cuda")
            L.check(lib.ts_per_sample_p2p_owned(
                L.ptr(d["tree"]), cap, L.ptr(ud), bsz, L.ptr(d["slots"]), L.ptr(peers),
                L.ptr(d["state"]), G, r, L.ptr(d["prio"]), 0.4, 0, L.ptr(idx), L.ptr(w),
                L.ptr(d["owned"]), L.ptr(d["err"]), L.stream()), "sample_p2p_owned")
            idxs.append(idx)
            mine = torch.nonzero(idx >= 0).reshape(-1)
            n = int(d["owned"][2])
            assert n == mine.numel(), (r, step, n, mine.numel())
            listed = d["owned"][4:4 + n].to(torch.int64).sort().values
            assert torch.equal(listed, mine), (r, step)
        got = sum(torch.where(i >= 0, i + r * cap, torch.zeros_like(i))
                  for r, i in enumerate(idxs)).cpu().numpy()
        assert np.array_equal(got, want), (G, step)
        glob.update_weight(want, td)
        for r, d in enumerate(ranks):
            # the plain update on a copy of the shard: scans all bsz entries
            ref_tree, ref_prio = d["tree"].clone(), d["prio"].clone()
            L.check(lib.ts_per_update(L.ptr(ref_tree), cap, cap, L.ptr(idxs[r]), L.ptr(tdd),
                                      L.TS_F32, bsz, 0.6, L.ptr(ref_prio), L.ptr(d["stamp"]),
                                      L.ptr(d["counter"]), L.ptr(d["err"]), L.stream()), "update")
            L.check(lib.ts_per_update_publish_owned(
                L.ptr(d["tree"]), cap, cap, L.ptr(idxs[r]), L.ptr(tdd), L.TS_F32, bsz, 0.6,
                L.ptr(d["prio"]), L.ptr(d["stamp"]), L.ptr(d["counter"]), L.ptr(d["err"]),
                L.ptr(d["state"]), L.ptr(peers), G, r, L.ptr(d["owned"]), L.stream()),
                "update_publish_owned")
            assert torch.equal(d["tree"], ref_tree), (r, step)
            assert torch.equal(d["prio"], ref_prio), (r, step)
            assert int(d["err"].item()) == 0
            # re-anchor the touched leaves on the oracle's f32 powers, then publish again
            loc = idxs[r][idxs[r] >= 0].unique()
            if loc.numel():
                vals = torch.from_numpy(glob.weight[loc.cpu().numpy() + r * cap]).cuda()
                L.check(lib.ts_segtree_update(L.ptr(d["tree"]), cap, cap, L.ptr(loc), L.ptr(vals),
                                              L.TS_F64, 1, loc.numel(), L.ptr(d["stamp"]),
                                              L.ptr(d["counter"]), L.ptr(d["err"]), L.stream()),
                        "anchor")
        for r, d in enumerate(ranks):
            L.check(lib.ts_per_publish(L.ptr(d["tree"]), L.ptr(d["prio"]), L.ptr(d["state"]),
                                       L.ptr(peers), G, r, L.stream()), "publish")


@pytest.mark.parametrize("G,cap,bsz", [(2, 1 << 13, 700), (4, 1 << 14, 4096)])
def test_p2p_weight_maximum_exchange_on_concurrent_streams(G, cap, bsz):
    """The MUTUAL half of the in-kernel exchange (every rank stores its weight maximum into every
    peer's slot buffer and waits for theirs) on one GPU: the G ranks' sampling kernels run
    CONCURRENTLY on G streams (G clusters of 16 CTAs are co-resident on a B200), so each one's
    bounded wait is answered by the others.  Normalised weights must equal the oracle's
    w / max(w) over the whole batch (prio.py:101); a peer timeout would raise the error flag."""
    from tianshou_fork_b200 import _lib as L
    lib = L.lib()
    rng = np.random.default_rng(G * 13 + 2)
    total = G * cap
    leaves = (np.abs(rng.standard_normal(total)) + 1e-7) ** 0.6
    glob = orc.PER(total, 0.6, 0.4)
    glob.weight[np.arange(total)] = leaves
    ranks = []
    for r in range(G):
        d = dict(tree=torch.zeros(2 * cap, dtype=torch.float64, device="cuda"),
                 stamp=torch.zeros(cap, dtype=torch.int32, device="cuda"),
                 counter=torch.zeros(4, dtype=torch.int32, device="cuda"),
                 err=torch.zeros(1, dtype=torch.int32, device="cuda"),
                 prio=torch.ones(2, dtype=torch.float64, device="cuda"),
                 slots=torch.zeros(16 * G, dtype=torch.float64, device="cuda"),
                 state=torch.zeros(2, dtype=torch.int64, device="cuda"),
                 idx=torch.empty(bsz, dtype=torch.int64, device="cuda"),
                 w=torch.empty(bsz, dtype=torch.float64, device="cuda"),
                 stream=torch.cuda.Stream())
        _fill(d["tree"], d["stamp"], d["counter"], d["err"], leaves[r * cap:(r + 1) * cap], L)
        ranks.append(d)
    peers = torch.tensor([d["slots"].data_ptr() for d in ranks], dtype=torch.int64, device="cuda")
    for r, d in enumerate(ranks):
        L.check(lib.ts_per_publish(L.ptr(d["tree"]), L.ptr(d["prio"]), L.ptr(d["state"]),
                                   L.ptr(peers), G, r, L.stream()), "publish")
    torch.cuda.synchronize()
    for step in range(3):
        u = rng.random(bsz)
        want = glob.sample_indices(u)
        want_w = glob.batch_weight(want)
        ud = torch.from_numpy(u).cuda()
        torch.cuda.synchronize()
        for r, d in enumerate(ranks):  # no dependencies between the streams: the kernels overlap
            L.check(lib.ts_per_sample_p2p(
                L.ptr(d["tree"]), cap, L.ptr(ud), bsz, L.ptr(d["slots"]), L.ptr(peers),
                L.ptr(d["state"]), G, r, L.ptr(d["prio"]), 0.4, 1, L.ptr(d["idx"]), L.ptr(d["w"]),
                L.ptr(d["err"]), d["stream"].cuda_stream), "sample_p2p")
        torch.cuda.synchronize()
        for r, d in enumerate(ranks):
            assert int(d["err"].item()) == 0, (r, step, "peer timeout or range flag")
        got = sum(torch.where(d["idx"] >= 0, d["idx"] + r * cap, torch.zeros_like(d["idx"]))
                  for r, d in enumerate(ranks)).cpu().numpy()
        assert np.array_equal(got, want), (G, step)
        got_w = sum(d["w"] for d in ranks).cpu().numpy()  # non-owners hold 0
        assert np.allclose(got_w, want_w, rtol=1e-12, atol=0), (G, step)
        assert float(got_w.max()) == 1.0
