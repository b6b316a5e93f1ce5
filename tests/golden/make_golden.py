"""Generate golden fixtures by running the REFERENCE itself (``/root/reference``).

Run in the build container only (the GPU box has no ``/root/reference``):

    python tests/golden/make_golden.py

The reference hard-imports ``gymnasium``, ``h5py``, ``pettingzoo`` and ``jsonargparse``,
none of which does arithmetic on the hot path and none of which is installed here; they
are replaced by attribute-only stub modules.  Everything written below is
an OUTPUT of reference code: ``BasePolicy.compute_episodic_return``,
``BasePolicy.compute_nstep_return``, ``ReplayBuffer`` / ``VectorReplayBuffer`` /
``ReplayBufferManager`` ``prev/next/unfinished_index/sample_indices``, ``SegmentTree`` and
``PrioritizedVectorReplayBuffer``.  Versions are recorded in ``versions.json``.
"""
from __future__ import annotations

import json
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("TS_REFERENCE", "/root/reference")


def import_reference():
    class _Any:
        def __init__(self, *a, **k):
            pass

        def __getattr__(self, k):
            return _Any()

        def __call__(self, *a, **k):
            return _Any()

    def stub(name, **attrs):
        m = types.ModuleType(name)
        m.__dict__.update(attrs)

        def _getattr(k):
            if k == "__version__":
                return "99.0.0"
            if k.startswith("__"):
                raise AttributeError(k)
            return type(k, (_Any,), {})

        m.__getattr__ = _getattr  # type: ignore
        sys.modules[name] = m
        return m

    for name in [
        "gymnasium", "gymnasium.spaces", "h5py", "pettingzoo", "pettingzoo.utils",
        "pettingzoo.utils.env", "pettingzoo.utils.wrappers", "jsonargparse",
    ]:
        if name not in sys.modules:
            stub(name)
    sys.modules["gymnasium"].spaces = sys.modules["gymnasium.spaces"]
    sys.path.insert(0, REF)
    import torch  # noqa: F401  (before the stubs are consulted by anything else)
    import tianshou  # noqa

    return tianshou


def fill_vecbuf(buf, Batch, rng, steps, n_env, p_term, p_trunc, ragged=False):
    """Drive ``buf.add`` like the Collector does (collector.py:324)."""
    for _ in range(steps):
        ids = np.arange(n_env)
        if ragged:
            ids = ids[rng.random(n_env) < 0.7]
            if ids.size == 0:
                continue
        m = ids.size
        term = rng.random(m) < p_term
        trunc = (rng.random(m) < p_trunc) & ~term
        batch = Batch(
            obs=rng.standard_normal((m, 3)).astype(np.float32),
            act=rng.integers(0, 4, m),
            rew=rng.standard_normal(m),
            terminated=term,
            truncated=trunc,
            obs_next=rng.standard_normal((m, 3)).astype(np.float32),
            info={},
        )
        buf.add(batch, buffer_ids=ids)


def gen_gae(ts, out):
    from tianshou.data import Batch, VectorReplayBuffer
    from tianshou.policy import BasePolicy

    cases = {}
    specs = [
        # name, n_env, steps, capacity_per_env, gamma, lam, p_term, p_trunc, ragged
        ("small", 4, 37, 64, 0.99, 0.95, 0.1, 0.05, False),
        ("wrap", 3, 150, 100, 0.99, 0.95, 0.05, 0.02, False),  # ring wrapped around
        ("ragged", 5, 60, 80, 0.9, 0.8, 0.1, 0.05, True),
        ("long4096", 2, 4096, 4096, 0.99, 0.95, 1 / 500, 1 / 2000, False),
        ("mc4096", 2, 4096, 4096, 1.0, 1.0, 1 / 1000, 0.0, False),  # gamma=lambda=1 stress
        ("noend", 2, 512, 512, 0.99, 0.95, 0.0, 0.0, False),
    ]
    for i, (name, n_env, steps, cap, gamma, lam, pt, ptr, ragged) in enumerate(specs):
        rng = np.random.default_rng(100 + i)
        buf = VectorReplayBuffer(cap * n_env, n_env)
        fill_vecbuf(buf, Batch, rng, steps, n_env, pt, ptr, ragged)
        batch, indices = buf.sample(0)
        n = len(indices)
        v_s = rng.standard_normal(n).astype(np.float32)
        v_s_ = rng.standard_normal(n).astype(np.float32)
        ret, adv = BasePolicy.compute_episodic_return(
            batch, buf, indices, v_s_=v_s_.copy(), v_s=v_s.copy(), gamma=gamma, gae_lambda=lam
        )
        cases[name] = dict(
            n_env=n_env, cap=cap, gamma=gamma, lam=lam,
            buf_rew=buf.rew.copy(), buf_terminated=buf.terminated.copy(),
            buf_truncated=buf.truncated.copy(), buf_done=buf.done.copy(),
            last_index=buf.last_index.copy(), lengths=buf._lengths.copy(),
            ring_index=np.array([b._index for b in buf.buffers]),
            indices=indices, unfinished=buf.unfinished_index(),
            v_s=v_s, v_s_=v_s_, returns=ret, adv=adv,
        )
        # v_s=None -> np.roll branch (policy/base.py:377); only for one small case
        if name == "small":
            ret2, adv2 = BasePolicy.compute_episodic_return(
                batch, buf, indices, v_s_=v_s_.copy(), v_s=None, gamma=gamma, gae_lambda=lam
            )
            cases[name].update(returns_roll=ret2, adv_roll=adv2)
            # v_s_=None requires lambda == 1 (:371-373), as PGPolicy does (pg.py:108-112)
            ret3, adv3 = BasePolicy.compute_episodic_return(
                batch, buf, indices, v_s_=None, v_s=None, gamma=gamma, gae_lambda=1.0
            )
            cases[name].update(returns_mc=ret3, adv_mc=adv3)
    flat = {f"{c}/{k}": np.asarray(v) for c, d in cases.items() for k, v in d.items()}
    np.savez_compressed(os.path.join(out, "gae.npz"), **flat)


def gen_index(ts, out):
    from tianshou.data import Batch, ReplayBuffer, ReplayBufferManager, VectorReplayBuffer

    flat = {}
    # equal sub-buffers, several fill levels incl. wrap-around and empty sub-buffers
    specs = [("v_partial", 4, 20, 11, True), ("v_wrap", 4, 20, 57, False),
             ("v_big", 16, 1024, 700, True), ("v_one", 1, 16, 40, False)]
    for i, (name, n_env, total, steps, ragged) in enumerate(specs):
        rng = np.random.default_rng(200 + i)
        buf = VectorReplayBuffer(total, n_env)
        fill_vecbuf(buf, Batch, rng, steps, n_env, 0.15, 0.05, ragged)
        q = np.concatenate([np.arange(-3, buf.maxsize + 3), rng.integers(0, buf.maxsize, 64)])
        flat.update({
            f"{name}/offset": buf._extend_offset, f"{name}/done": buf.done,
            f"{name}/last_index": buf.last_index, f"{name}/lengths": buf._lengths,
            f"{name}/query": q, f"{name}/prev": buf.prev(q), f"{name}/next": buf.next(q),
            f"{name}/unfinished": buf.unfinished_index(),
            f"{name}/all_indices": buf.sample_indices(0),
            f"{name}/ring_index": np.array([b._index for b in buf.buffers]),
        })
    # unequal sub-buffers (plain ReplayBufferManager)
    rng = np.random.default_rng(250)
    buf = ReplayBufferManager([ReplayBuffer(5), ReplayBuffer(9), ReplayBuffer(3), ReplayBuffer(16)])
    fill_vecbuf(buf, Batch, rng, 12, 4, 0.2, 0.1, True)
    q = np.arange(-2, buf.maxsize + 2)
    flat.update({
        "m_unequal/offset": buf._extend_offset, "m_unequal/done": buf.done,
        "m_unequal/last_index": buf.last_index, "m_unequal/lengths": buf._lengths,
        "m_unequal/query": q, "m_unequal/prev": buf.prev(q), "m_unequal/next": buf.next(q),
        "m_unequal/unfinished": buf.unfinished_index(),
        "m_unequal/all_indices": buf.sample_indices(0),
        "m_unequal/ring_index": np.array([b._index for b in buf.buffers]),
    })
    # single ring, partially filled then wrapped
    for name, size, steps in [("s_partial", 16, 9), ("s_wrap", 16, 41)]:
        rng = np.random.default_rng(260 + steps)
        buf = ReplayBuffer(size)
        for _ in range(steps):
            term = rng.random() < 0.2
            buf.add(Batch(obs=np.zeros(2), act=0, rew=rng.standard_normal(), terminated=term,
                          truncated=(not term) and rng.random() < 0.1, obs_next=np.zeros(2),
                          info={}))
        q = np.arange(len(buf))
        flat.update({
            f"{name}/done": buf.done, f"{name}/last_index": buf.last_index,
            f"{name}/size": np.array(len(buf)), f"{name}/ring_index": np.array(buf._index),
            f"{name}/query": q, f"{name}/prev": buf.prev(q), f"{name}/next": buf.next(q),
            f"{name}/unfinished": buf.unfinished_index(),
            f"{name}/all_indices": buf.sample_indices(0),
        })
    np.savez_compressed(os.path.join(out, "index.npz"), **flat)


def gen_nstep(ts, out):
    import torch
    from tianshou.data import Batch, VectorReplayBuffer
    from tianshou.policy import BasePolicy

    flat = {}
    specs = [("n1", 4, 64, 40, 1, 1), ("n3", 4, 64, 100, 3, 1), ("n3_d51", 4, 64, 100, 3, 51),
             ("n10", 8, 256, 300, 10, 1), ("n5_partial", 6, 600, 37, 5, 3)]
    for i, (name, n_env, total, steps, n_step, D) in enumerate(specs):
        rng = np.random.default_rng(300 + i)
        buf = VectorReplayBuffer(total, n_env)
        fill_vecbuf(buf, Batch, rng, steps, n_env, 0.1, 0.05, name == "n5_partial")
        table = rng.standard_normal((buf.maxsize, D)).astype(np.float32)

        def target_q_fn(buffer, indices, table=table, D=D):
            t = torch.from_numpy(table[indices])
            return t if D > 1 else t.flatten()

        indice = buf.sample_indices(0)
        indice = np.concatenate([indice, rng.choice(indice, 50)])
        batch = Batch()
        out_b = BasePolicy.compute_nstep_return(batch, buf, indice, target_q_fn, gamma=0.97,
                                                n_step=n_step)
        flat.update({
            f"{name}/n_env": np.array(n_env), f"{name}/n_step": np.array(n_step),
            f"{name}/gamma": np.array(0.97), f"{name}/table": table,
            f"{name}/rew": buf.rew, f"{name}/done": buf.done,
            f"{name}/terminated": buf.terminated, f"{name}/truncated": buf.truncated,
            f"{name}/offset": buf._extend_offset, f"{name}/last_index": buf.last_index,
            f"{name}/lengths": buf._lengths, f"{name}/unfinished": buf.unfinished_index(),
            f"{name}/ring_index": np.array([b._index for b in buf.buffers]),
            f"{name}/indice": indice, f"{name}/returns": out_b.returns.numpy(),
        })
    np.savez_compressed(os.path.join(out, "nstep.npz"), **flat)


def gen_segtree(ts, out):
    from tianshou.data import SegmentTree

    flat = {}
    for name, size, rounds, bsz in [("t8", 8, 30, 4), ("t100", 100, 40, 16),
                                    ("t16384", 16384, 20, 256), ("t5000", 5000, 10, 1024)]:
        rng = np.random.default_rng(400 + size)
        tree = SegmentTree(size)
        idx_log, val_log = [], []
        for r in range(rounds):
            idx = rng.integers(0, size, bsz)
            if r % 3 == 0:  # force duplicates: last occurrence must win
                idx[bsz // 2:] = idx[: bsz - bsz // 2]
            val = rng.random(bsz) if r % 2 == 0 else rng.random(bsz).astype(np.float32)
            tree[idx] = val
            idx_log.append(idx)
            val_log.append(val.astype(np.float64))
        total = tree.reduce()
        u = rng.random(512)
        scalar = u * total
        pidx = tree.get_prefix_sum_idx(scalar.copy())
        scalar32 = (u * total * 0.999).astype(np.float32)
        mutated32 = scalar32.copy()
        pidx32 = tree.get_prefix_sum_idx(mutated32)
        mutated = scalar.copy()
        tree.get_prefix_sum_idx(mutated)
        se = np.sort(rng.integers(0, size + 1, (64, 2)), axis=1)
        se = se[se[:, 0] < se[:, 1]]
        red = np.array([tree.reduce(int(s), int(e)) for s, e in se])
        flat.update({
            f"{name}/size": np.array(size), f"{name}/idx_log": np.stack(idx_log),
            f"{name}/val_log": np.stack(val_log), f"{name}/tree": tree._value.copy(),
            f"{name}/scalar": scalar, f"{name}/prefix_idx": pidx, f"{name}/scalar_after": mutated,
            f"{name}/scalar32": scalar32, f"{name}/prefix_idx32": pidx32,
            f"{name}/scalar32_after": mutated32,
            f"{name}/reduce_se": se, f"{name}/reduce": red,
            f"{name}/reduce_neg_end": np.array(tree.reduce(1, -1)),
        })
    np.savez_compressed(os.path.join(out, "segtree.npz"), **flat)


def gen_per(ts, out):
    from tianshou.data import Batch, PrioritizedReplayBuffer, PrioritizedVectorReplayBuffer

    flat = {}
    for name, n_env, total, steps, bsz, wn in [("pv", 4, 256, 90, 32, True),
                                                ("pv_nonorm", 8, 4096, 600, 256, False)]:
        rng = np.random.default_rng(500 + n_env)
        buf = PrioritizedVectorReplayBuffer(total, n_env, alpha=0.6, beta=0.4, weight_norm=wn)
        fill_vecbuf(buf, Batch, rng, steps, n_env, 0.1, 0.05, False)
        flat[f"{name}/tree_after_add"] = buf.weight._value.copy()
        trees, idxs, weights, tds, us, mx, mn = [], [], [], [], [], [], []
        for it in range(6):
            np.random.seed(1000 + it)
            u = np.random.rand(bsz)
            np.random.seed(1000 + it)
            idx = buf.sample_indices(bsz)
            w = buf[idx].weight
            td = rng.standard_normal(bsz).astype(np.float32)
            if it % 2 == 0:
                idx[-4:] = idx[:4]  # duplicates inside update_weight
            buf.update_weight(idx, td)
            us.append(u); idxs.append(idx.copy()); weights.append(np.asarray(w)); tds.append(td)
            trees.append(buf.weight._value.copy())
            mx.append(float(buf._max_prio)); mn.append(float(buf._min_prio))
        flat.update({
            f"{name}/size": np.array(buf.maxsize), f"{name}/n_env": np.array(n_env),
            f"{name}/u": np.stack(us), f"{name}/idx": np.stack(idxs),
            f"{name}/weight": np.stack(weights), f"{name}/td": np.stack(tds),
            f"{name}/tree": np.stack(trees), f"{name}/max_prio": np.array(mx),
            f"{name}/min_prio": np.array(mn), f"{name}/lengths": buf._lengths,
        })
    # single prioritized ring: test_priortized_replaybuffer-style (test_buffer.py:256-302)
    rng = np.random.default_rng(550)
    buf = PrioritizedReplayBuffer(50, 0.5, 0.5)
    for i in range(70):
        buf.add(Batch(obs=np.zeros(2), act=0, rew=1.0, terminated=i % 7 == 6, truncated=False,
                      obs_next=np.zeros(2), info={}))
    idx = rng.integers(0, 50, 20)
    td = rng.standard_normal(20)
    buf.update_weight(idx, td)
    flat.update({"single/idx": idx, "single/td": td, "single/tree": buf.weight._value.copy(),
                 "single/max_prio": np.array(buf._max_prio),
                 "single/min_prio": np.array(buf._min_prio),
                 "single/weight_all": buf.get_weight(np.arange(50))})
    np.savez_compressed(os.path.join(out, "per.npz"), **flat)


def gen_ppo(ts, out):
    """Loss values and gradients of the reference expressions at ppo.py:109-144 evaluated
    by torch autograd in f64 on CPU (the reference pins no loss value in its tests)."""
    import torch

    flat = {}
    rng = np.random.default_rng(600)
    n = 1000
    base = dict(
        logp_old=rng.normal(-8, 0.3, n), adv=rng.standard_normal(n) * 3 + 0.5,
        returns=rng.standard_normal(n) * 2, v_s=rng.standard_normal(n),
    )
    base["logp"] = base["logp_old"] + rng.normal(0, 0.25, n)
    base["value"] = base["v_s"] + rng.normal(0, 0.3, n)
    for k, v in base.items():
        flat[f"in/{k}"] = v.astype(np.float32)
    variants = [("plain", False, False, None), ("norm", True, False, None),
                ("vclip", False, True, None), ("norm_vclip", True, True, None),
                ("dual", True, True, 5.0), ("dual_nonorm", False, False, 1.5)]
    eps_clip, eps = 0.2, 1e-8
    for name, norm_adv, value_clip, dual_clip in variants:
        t = {k: torch.tensor(v.astype(np.float32).astype(np.float64)) for k, v in base.items()}
        logp = t["logp"].clone().requires_grad_(True)
        value = t["value"].clone().requires_grad_(True)
        adv = t["adv"]
        if norm_adv:
            mean, std = adv.mean(), adv.std()
            adv = (adv - mean) / (std + eps)
        ratio = (logp - t["logp_old"]).exp()
        surr1 = ratio * adv
        surr2 = ratio.clamp(1.0 - eps_clip, 1.0 + eps_clip) * adv
        if dual_clip:
            clip1 = torch.min(surr1, surr2)
            clip2 = torch.max(clip1, dual_clip * adv)
            clip_loss = -torch.where(adv < 0, clip2, clip1).mean()
        else:
            clip_loss = -torch.min(surr1, surr2).mean()
        if value_clip:
            v_clip = t["v_s"] + (value - t["v_s"]).clamp(-eps_clip, eps_clip)
            vf1 = (t["returns"] - value).pow(2)
            vf2 = (t["returns"] - v_clip).pow(2)
            vf_loss = torch.max(vf1, vf2).mean()
        else:
            vf_loss = (t["returns"] - value).pow(2).mean()
        (clip_loss + 0.5 * vf_loss).backward()
        flat.update({
            f"{name}/cfg": np.array([float(norm_adv), float(value_clip), dual_clip or 0.0,
                                     eps_clip, 0.5]),
            f"{name}/clip_loss": clip_loss.detach().numpy(),
            f"{name}/vf_loss": vf_loss.detach().numpy(),
            f"{name}/dlogp": logp.grad.numpy(), f"{name}/dvalue": value.grad.numpy(),
            f"{name}/adv_used": adv.numpy(),
        })
    np.savez_compressed(os.path.join(out, "ppo.npz"), **flat)


PPO_UPDATE_SPECS = [
    # name, kwargs of PPOPolicy, batch_size, repeat, number of update() calls
    ("clip_norm", dict(eps_clip=0.2, value_clip=True, advantage_normalization=True, vf_coef=0.5,
                       ent_coef=0.01, max_grad_norm=0.5, gae_lambda=0.95, discount_factor=0.99),
     64, 2, 1),
    ("rewnorm_recompute_dual", dict(eps_clip=0.1, dual_clip=3.0, value_clip=False,
                                    advantage_normalization=False, recompute_advantage=True,
                                    reward_normalization=True, vf_coef=0.25, ent_coef=0.0,
                                    max_grad_norm=None, gae_lambda=0.9, discount_factor=0.97,
                                    max_batchsize=300),
     128, 3, 2),
]


def ppo_update_nets(torch):
    """Actor / critic of the PPO-update goldens (shared with tests/test_gpu_ppo.py)."""

    class Actor(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.net = torch.nn.Sequential(torch.nn.Linear(3, 64), torch.nn.Tanh(),
                                           torch.nn.Linear(64, 4))

        def forward(self, obs, state=None, info=None):
            obs = torch.as_tensor(obs, dtype=torch.float32,
                                  device=self.net[0].weight.device)
            return self.net(obs), state

    class Critic(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.net = torch.nn.Sequential(torch.nn.Linear(3, 64), torch.nn.Tanh(),
                                           torch.nn.Linear(64, 1))

        def forward(self, obs, **kw):
            obs = torch.as_tensor(obs, dtype=torch.float32,
                                  device=self.net[0].weight.device)
            return self.net(obs)

    return Actor, Critic


def gen_ppo_update(ts, out):
    """``PPOPolicy.update`` of the reference itself on CPU (policy/base.py:288-315 ->
    ppo.py:87-162, a2c.py:83-117) over a 20 x 100 rollout (test/discrete/test_ppo.py:29-33
    shape): returns / advantages / logp_old of the first ``process_fn``, the logged losses
    of every minibatch, the return statistics and the parameters after the update(s)."""
    import torch
    from tianshou.data import Batch, VectorReplayBuffer
    from tianshou.policy import PPOPolicy

    Actor, Critic = ppo_update_nets(torch)
    flat = {}
    for name, kw, batch_size, repeat, n_updates in PPO_UPDATE_SPECS:
        buf = VectorReplayBuffer(20 * 100, 20)
        fill_vecbuf(buf, Batch, np.random.default_rng(700), 100, 20, 0.03, 0.01)
        torch.manual_seed(1)
        actor, critic = Actor(), Critic()
        optim = torch.optim.Adam(list(actor.parameters()) + list(critic.parameters()), lr=3e-4)
        for k, v in actor.state_dict().items():
            flat[f"{name}/init/actor/{k}"] = v.numpy().copy()
        for k, v in critic.state_dict().items():
            flat[f"{name}/init/critic/{k}"] = v.numpy().copy()
        policy = PPOPolicy(actor, critic, optim,
                           lambda logits: torch.distributions.Categorical(logits=logits),
                           action_scaling=False, **kw)
        seen = []
        inner = policy.process_fn

        def spy(batch, buffer, indices, inner=inner, seen=seen):
            batch = inner(batch, buffer, indices)
            seen.append({k: np.asarray(batch[k].detach().numpy()).copy()
                         for k in ("returns", "adv", "v_s", "logp_old")})
            return batch

        policy.process_fn = spy
        np.random.seed(7)
        for u in range(n_updates):
            res = policy.update(0, buf, batch_size=batch_size, repeat=repeat)
            for k, v in res.items():
                flat[f"{name}/u{u}/{k}"] = np.asarray(v, dtype=np.float64)
            for k, v in seen[u].items():
                flat[f"{name}/u{u}/{k}"] = v
            flat[f"{name}/u{u}/ret_rms"] = np.array(
                [policy.ret_rms.mean, policy.ret_rms.var, policy.ret_rms.count], np.float64)
        for k, v in actor.state_dict().items():
            flat[f"{name}/final/actor/{k}"] = v.numpy().copy()
        for k, v in critic.state_dict().items():
            flat[f"{name}/final/critic/{k}"] = v.numpy().copy()
    np.savez_compressed(os.path.join(out, "ppo_update.npz"), **flat)


def gen_pg_a2c_update(ts, out):
    """``PGPolicy.update`` (pg.py:87-120, 174-192) and ``A2CPolicy.update`` (a2c.py:77-161) of the
    reference on the rollout of gen_ppo_update, with return normalisation on, two updates."""
    import torch
    from tianshou.data import Batch, VectorReplayBuffer
    from tianshou.policy import A2CPolicy, PGPolicy

    Actor, Critic = ppo_update_nets(torch)
    flat = {}
    for name in ("pg", "a2c"):
        buf = VectorReplayBuffer(20 * 100, 20)
        fill_vecbuf(buf, Batch, np.random.default_rng(700), 100, 20, 0.03, 0.01)
        torch.manual_seed(2)
        actor, critic = Actor(), Critic()
        dist_fn = lambda logits: torch.distributions.Categorical(logits=logits)  # noqa: E731
        if name == "pg":
            optim = torch.optim.Adam(actor.parameters(), lr=1e-3)
            policy = PGPolicy(actor, optim, dist_fn, discount_factor=0.95,
                              reward_normalization=True, action_scaling=False)
            keys = ("returns",)
        else:
            optim = torch.optim.Adam(list(actor.parameters()) + list(critic.parameters()),
                                     lr=1e-3)
            policy = A2CPolicy(actor, critic, optim, dist_fn, vf_coef=0.5, ent_coef=0.01,
                               max_grad_norm=0.5, gae_lambda=0.9, discount_factor=0.98,
                               reward_normalization=True, action_scaling=False)
            keys = ("returns", "adv", "v_s")
        for tag, net in (("actor", actor), ("critic", critic)):
            for k, v in net.state_dict().items():
                flat[f"{name}/init/{tag}/{k}"] = v.numpy().copy()
        seen = []
        inner = policy.process_fn

        def spy(batch, buffer, indices, inner=inner, seen=seen, keys=keys):
            batch = inner(batch, buffer, indices)
            seen.append({k: np.asarray(batch[k].detach().numpy() if hasattr(batch[k], "detach")
                                       else batch[k]).copy() for k in keys})
            return batch

        policy.process_fn = spy
        np.random.seed(9)
        for u in range(2):
            res = policy.update(0, buf, batch_size=100, repeat=2)
            for k, v in res.items():
                flat[f"{name}/u{u}/{k}"] = np.asarray(v, dtype=np.float64)
            for k, v in seen[u].items():
                flat[f"{name}/u{u}/{k}"] = v
            flat[f"{name}/u{u}/ret_rms"] = np.array(
                [policy.ret_rms.mean, policy.ret_rms.var, policy.ret_rms.count], np.float64)
        for tag, net in (("actor", actor), ("critic", critic)):
            for k, v in net.state_dict().items():
                flat[f"{name}/final/{tag}/{k}"] = v.numpy().copy()
    np.savez_compressed(os.path.join(out, "pg_a2c_update.npz"), **flat)


DQN_UPDATE_SPECS = [
    # name, DQNPolicy kwargs, PER kwargs, sample size, updates
    ("double_n3", dict(discount_factor=0.97, estimation_step=3, target_update_freq=2,
                       is_double=True), dict(alpha=0.6, beta=0.4), 64, 6),
    ("nature_n1_huber", dict(discount_factor=0.9, estimation_step=1, target_update_freq=0,
                             is_double=False, clip_loss_grad=True),
     dict(alpha=0.8, beta=1.0), 48, 4),
]


def gen_dqn_update(ts, out):
    """``DQNPolicy.update`` of the reference (dqn.py:98-194 through policy/base.py:288-315) on a
    ``PrioritizedVectorReplayBuffer``: per update the sampled indices, the importance weights
    handed to ``process_fn``, the n-step returns, the loss, and after ``post_process_fn`` the
    leaves at the sampled indices plus ``_max_prio`` / ``_min_prio``; parameters at the end."""
    import torch
    from tianshou.data import Batch, PrioritizedVectorReplayBuffer
    from tianshou.policy import DQNPolicy

    Actor, _ = ppo_update_nets(torch)
    flat = {}
    for name, kw, per_kw, bsz, n_updates in DQN_UPDATE_SPECS:
        buf = PrioritizedVectorReplayBuffer(20 * 100, 20, **per_kw)
        fill_vecbuf(buf, Batch, np.random.default_rng(800), 100, 20, 0.03, 0.01)
        torch.manual_seed(3)
        net = Actor()
        optim = torch.optim.Adam(net.parameters(), lr=1e-3)
        for k, v in net.state_dict().items():
            flat[f"{name}/init/actor/{k}"] = v.numpy().copy()
        policy = DQNPolicy(net, optim, **kw)
        seen = []
        inner = policy.process_fn

        def spy(batch, buffer, indices, inner=inner, seen=seen):
            rec = {"indices": np.asarray(indices).copy(),
                   "weight_in": np.asarray(batch.weight).copy()}
            batch = inner(batch, buffer, indices)
            rec["returns"] = batch.returns.detach().numpy().copy()
            seen.append(rec)
            return batch

        policy.process_fn = spy
        np.random.seed(13)
        for u in range(n_updates):
            res = policy.update(bsz, buf)
            for k, v in seen[u].items():
                flat[f"{name}/u{u}/{k}"] = v
            flat[f"{name}/u{u}/loss"] = np.float64(res["loss"])
            flat[f"{name}/u{u}/leaves"] = np.asarray(buf.weight[seen[u]["indices"]], np.float64)
            flat[f"{name}/u{u}/prio"] = np.array([buf._max_prio, buf._min_prio], np.float64)
        for k, v in net.state_dict().items():
            flat[f"{name}/final/actor/{k}"] = v.numpy().copy()
    np.savez_compressed(os.path.join(out, "dqn_update.npz"), **flat)


def gen_sample(ts, out):
    """Seeded ``sample_indices(batch_size > 0)`` of the reference (base.py:276-305,
    manager.py:163-192): the draws come from the global ``np.random`` stream, so a port that
    makes the same calls in the same order returns the same indices."""
    from tianshou.data import Batch, ReplayBuffer, VectorReplayBuffer

    flat = {}
    # name, buffers, total, steps, kwargs
    specs = [("vec_partial", 4, 40, 7, {}), ("vec_wrapped", 3, 30, 25, {}),
             ("vec_stack_avail", 4, 64, 11, dict(stack_num=3, sample_avail=True)),
             ("vec_one_empty", 5, 50, 6, {})]
    for name, nb, total, steps, kw in specs:
        buf = VectorReplayBuffer(total, nb, **kw)
        rng = np.random.default_rng(950)
        for t in range(steps):
            ids = np.arange(nb if name != "vec_one_empty" else nb - 1)
            m = ids.size
            term = rng.random(m) < 0.15
            buf.add(Batch(obs=rng.standard_normal((m, 2)).astype(np.float32),
                          act=rng.integers(0, 3, m), rew=rng.standard_normal(m),
                          terminated=term, truncated=np.zeros(m, bool),
                          obs_next=rng.standard_normal((m, 2)).astype(np.float32), info={}),
                    buffer_ids=ids)
        np.random.seed(31)
        for i, bsz in enumerate((1, 7, 64)):
            flat[f"{name}/draw{i}"] = buf.sample_indices(bsz)
    for name, size, steps, kw in [("ring", 16, 9, {}), ("ring_wrapped", 16, 41, {}),
                                  ("ring_stack_avail", 32, 20,
                                   dict(stack_num=4, sample_avail=True))]:
        buf = ReplayBuffer(size, **kw)
        rng = np.random.default_rng(951)
        for t in range(steps):
            buf.add(Batch(obs=rng.standard_normal(2).astype(np.float32), act=int(rng.integers(3)),
                          rew=float(rng.standard_normal()), terminated=bool(rng.random() < 0.2),
                          truncated=False, obs_next=rng.standard_normal(2).astype(np.float32),
                          info={}))
        np.random.seed(32)
        for i, bsz in enumerate((1, 5, 40)):
            flat[f"{name}/draw{i}"] = buf.sample_indices(bsz)
    np.savez_compressed(os.path.join(out, "sample.npz"), **flat)


STACK_SPECS = [
    # name, n_env, total, steps, stack_num, ragged, frame shape (make_golden.gen_stack)
    ("atari_like", 4, 80, 13, 4, False, (6,)),
    ("wrapped", 3, 30, 47, 4, True, (2, 3)),
    ("deep", 2, 64, 25, 7, False, (4,)),
]


def stack_stream(rng, steps, n_env, ragged, frame_shape, hist):
    """Add-stream of save_only_last_obs buffers: obs carries `hist` frames, the last is kept."""
    for _ in range(steps):
        ids = np.arange(n_env)
        if ragged:
            ids = ids[rng.random(n_env) < 0.7]
            if ids.size == 0:
                continue
        m = ids.size
        term = rng.random(m) < 0.2
        trunc = (rng.random(m) < 0.1) & ~term
        yield ids, dict(
            obs=rng.integers(0, 256, (m, hist) + frame_shape).astype(np.uint8),
            act=rng.integers(0, 4, m), rew=rng.standard_normal(m), terminated=term,
            truncated=trunc, info={})


def gen_stack(ts, out):
    """Frame-stack gathers (base.py:317-358, 374-378): buf[q].obs / .obs_next of
    ignore_obs_next + save_only_last_obs buffers (the Atari layout, atari_dqn.py:135-140),
    a single ring with saved obs_next, and get() with an explicit stack_num."""
    from tianshou.data import Batch, ReplayBuffer, VectorReplayBuffer

    flat = {}
    for i, (name, n_env, total, steps, k, ragged, fshape) in enumerate(STACK_SPECS):
        rng = np.random.default_rng(700 + i)
        buf = VectorReplayBuffer(total, n_env, stack_num=k, ignore_obs_next=True,
                                 save_only_last_obs=True)
        for ids, d in stack_stream(rng, steps, n_env, ragged, fshape, k):
            buf.add(Batch(**d), buffer_ids=ids)
        q = np.concatenate([np.arange(buf.maxsize), rng.integers(0, buf.maxsize, 40)])
        data = buf[q]
        flat.update({
            f"{name}/frames": buf.obs, f"{name}/offset": buf._extend_offset,
            f"{name}/done": buf.done, f"{name}/last_index": buf.last_index,
            f"{name}/lengths": buf._lengths, f"{name}/query": q, f"{name}/obs": data.obs,
            f"{name}/obs_next": data.obs_next,
            f"{name}/get2": buf.get(q, "obs", stack_num=2),
        })
    # single ring, float frames, obs_next saved (both keys go through get())
    rng = np.random.default_rng(750)
    buf = ReplayBuffer(16, stack_num=3)
    for _ in range(41):
        term = rng.random() < 0.2
        buf.add(Batch(obs=rng.standard_normal(5).astype(np.float32), act=0,
                      rew=rng.standard_normal(), terminated=term,
                      truncated=(not term) and rng.random() < 0.1,
                      obs_next=rng.standard_normal(5).astype(np.float32), info={}))
    q = np.arange(len(buf))
    data = buf[q]
    flat.update({"ring/frames": buf.obs, "ring/frames_next": buf.obs_next, "ring/done": buf.done,
                 "ring/last_index": buf.last_index, "ring/size": np.array(len(buf)),
                 "ring/query": q, "ring/obs": data.obs, "ring/obs_next": data.obs_next})
    np.savez_compressed(os.path.join(out, "stack.npz"), **flat)


def gen_c4(ts, out):
    """BASELINE.json configs[3] at FULL size on the reference itself:
    ``PrioritizedVectorReplayBuffer(2^20, 16)`` with wrapped rings, then for batch 32 / 256 /
    4096: ``sample_indices`` (prio.py:59-64), ``buffer[idx].weight`` (:91-102),
    ``compute_nstep_return`` n = 3 with D = 1 and 51 (policy/base.py:386-440),
    ``update_weight`` with 5 % duplicated indices (prio.py:78-89); plus ``prev`` / ``next`` of
    4096 random indices.  Inputs are regenerated from seeds by tests/helpers.py."""
    import hashlib

    import torch
    from tianshou.data import Batch, PrioritizedVectorReplayBuffer
    from tianshou.policy import BasePolicy

    sys.path.insert(0, os.path.dirname(HERE))
    import helpers as H

    rew, term, trunc = H.c4_stream()
    E, sub = H.C4_E, H.C4_SUB
    buf = PrioritizedVectorReplayBuffer(E * sub, E, alpha=0.6, beta=0.4)
    zeros = np.zeros(E, np.uint8)
    for t in range(rew.shape[0]):
        buf.add(Batch(obs=zeros, act=zeros, rew=rew[t], terminated=term[t], truncated=trunc[t],
                      obs_next=zeros, info={}))
    size = E * sub
    buf.weight[np.arange(size)] = H.c4_leaves()
    flat = {
        "c4/rew_sha": np.frombuffer(hashlib.sha256(buf.rew.tobytes()).digest(), np.uint8),
        "c4/done_sha": np.frombuffer(hashlib.sha256(buf.done.tobytes()).digest(), np.uint8),
        "c4/last_index": buf.last_index.copy(), "c4/lengths": buf._lengths.copy(),
        "c4/unfinished": buf.unfinished_index(), "c4/root0": np.array(buf.weight.reduce()),
    }
    rng = np.random.default_rng(4002)
    q = rng.integers(0, size, 4096)
    flat["c4/q"], flat["c4/prev"], flat["c4/next"] = q, buf.prev(q), buf.next(q)
    for bsz in H.C4_BATCHES:
        np.random.seed(2000 + bsz)
        idx = buf.sample_indices(bsz)
        batch = buf[idx]
        flat[f"c4/b{bsz}/idx"] = idx.copy()
        flat[f"c4/b{bsz}/weight"] = np.asarray(batch.weight).copy()
        for D in H.C4_DIMS:
            def target_q_fn(buffer, terminal, D=D):
                t = torch.from_numpy(H.c4_target_q(terminal, D))
                return t if D > 1 else t.flatten()

            r = BasePolicy.compute_nstep_return(Batch(), buf, idx, target_q_fn, gamma=0.99,
                                                n_step=3).returns.numpy()
            flat[f"c4/b{bsz}/d{D}/returns_sha"] = np.frombuffer(
                hashlib.sha256(np.ascontiguousarray(r).tobytes()).digest(), np.uint8)
            flat[f"c4/b{bsz}/d{D}/returns_head"] = r[:64].copy()
        td = rng.standard_normal(bsz).astype(np.float32)
        idx_u = idx.copy()
        k = bsz // 20
        idx_u[:k] = idx_u[k:2 * k]
        buf.update_weight(idx_u, td)
        uniq = np.unique(idx_u)
        flat[f"c4/b{bsz}/leaves"] = buf.weight[uniq].copy()
        flat[f"c4/b{bsz}/root"] = np.array(buf.weight.reduce())
        flat[f"c4/b{bsz}/prio"] = np.array([buf._max_prio, buf._min_prio])
    np.savez_compressed(os.path.join(out, "c4.npz"), **flat)


def gen_ppo_ant(ts, out):
    """BASELINE.json configs[2]: ``PPOPolicy.update`` of the reference on CPU over an Ant-shaped
    rollout (obs 27, act 8, 64 envs x 2048 steps = 131 072 transitions, Gaussian actor, minibatch
    1024, one repeat; policy/base.py:288-315 -> ppo.py:87-162, a2c.py:83-117).  Stored: a strided
    sample + sums of returns / adv / v_s / logp_old, every minibatch's losses, final weights."""
    import torch
    from tianshou.data import Batch, VectorReplayBuffer
    from tianshou.policy import PPOPolicy

    sys.path.insert(0, os.path.dirname(HERE))
    import helpers as H

    A = H.ANT
    obs, act, rew, term, trunc = H.ant_rollout()
    buf = VectorReplayBuffer(A["E"] * A["T"], A["E"])
    for t in range(A["T"]):
        buf.add(Batch(obs=obs[t], act=act[t], rew=rew[t], terminated=term[t], truncated=trunc[t],
                      obs_next=obs[t + 1], info={}))
    Actor, Critic = H.ant_nets(torch)
    torch.manual_seed(2)
    actor, critic = Actor(), Critic()
    optim = torch.optim.Adam(list(actor.parameters()) + list(critic.parameters()), lr=3e-4)
    flat = {}
    for k, v in actor.state_dict().items():
        flat[f"ant/init/actor/{k}"] = v.numpy().copy()
    for k, v in critic.state_dict().items():
        flat[f"ant/init/critic/{k}"] = v.numpy().copy()
    policy = PPOPolicy(actor, critic, optim, H.ant_dist(torch), action_scaling=False, **H.ANT_KW)
    seen = {}
    inner = policy.process_fn

    def spy(batch, buffer, indices):
        batch = inner(batch, buffer, indices)
        for k in ("returns", "adv", "v_s", "logp_old"):
            seen[k] = np.asarray(batch[k].detach().numpy(), dtype=np.float64).copy()
        return batch

    policy.process_fn = spy
    np.random.seed(13)
    res = policy.update(0, buf, batch_size=A["batch_size"], repeat=A["repeat"])
    for k, v in res.items():
        flat[f"ant/{k}"] = np.asarray(v, dtype=np.float64)
    for k, v in seen.items():
        flat[f"ant/{k}_sample"] = v[::37].astype(np.float32)
        flat[f"ant/{k}_stats"] = np.array([v.sum(), np.abs(v).sum(), (v * v).sum()])
    for k, v in actor.state_dict().items():
        flat[f"ant/final/actor/{k}"] = v.numpy().copy()
    for k, v in critic.state_dict().items():
        flat[f"ant/final/critic/{k}"] = v.numpy().copy()
    np.savez_compressed(os.path.join(out, "ppo_ant.npz"), **flat)


def main():
    ts = import_reference()
    import numba

    only = sys.argv[1:]  # e.g. `make_golden.py c4 ppo_ant` regenerates just those files
    if only:
        for name in only:
            globals()[f"gen_{name}"](ts, HERE)
        return

    gen_gae(ts, HERE)
    gen_index(ts, HERE)
    gen_nstep(ts, HERE)
    gen_segtree(ts, HERE)
    gen_per(ts, HERE)
    gen_ppo(ts, HERE)
    gen_ppo_update(ts, HERE)
    gen_pg_a2c_update(ts, HERE)
    gen_dqn_update(ts, HERE)
    gen_sample(ts, HERE)
    gen_stack(ts, HERE)
    gen_c4(ts, HERE)
    gen_ppo_ant(ts, HERE)
    with open(os.path.join(HERE, "versions.json"), "w") as f:
        json.dump({"tianshou": ts.__version__, "numpy": np.__version__,
                   "numba": numba.__version__, "reference": REF}, f, indent=1)
    for fn in sorted(os.listdir(HERE)):
        print(fn, os.path.getsize(os.path.join(HERE, fn)))


if __name__ == "__main__":
    main()
