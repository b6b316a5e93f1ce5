"""Fused PPO loss vs the reference expressions (golden: torch autograd in f64) and the oracle;
plus a CartPole-shaped end-to-end PPOPolicy.update smoke (BASELINE.json configs[0] shape)."""
import numpy as np
import pytest
import torch

from conftest import load_golden
from oracle import oracle as orc

pytestmark = pytest.mark.gpu

VARIANTS = ["plain", "norm", "vclip", "norm_vclip", "dual", "dual_nonorm"]


@pytest.mark.parametrize("name", VARIANTS)
def test_loss_and_gradients_vs_reference_expressions(name):
    import tianshou_fork_b200 as ts
    g = load_golden("ppo")
    x = {k: torch.from_numpy(v).cuda() for k, v in g["in"].items()}
    c = g[name]
    norm_adv, value_clip, dual, eps_clip, vf_coef = c["cfg"]
    logp = x["logp"].clone().requires_grad_(True)
    value = x["value"].clone().requires_grad_(True)
    clip_loss, vf_loss, stats = ts.ppo_loss(
        logp, value, x["logp_old"], x["adv"], x["returns"], x["v_s"], None, eps_clip,
        dual or None, bool(value_clip), bool(norm_adv))
    # loss parity <= 1e-5 relative vs the reference expression evaluated in f64
    assert np.isclose(clip_loss.item(), c["clip_loss"], rtol=1e-5)
    assert np.isclose(vf_loss.item(), c["vf_loss"], rtol=1e-5)
    (clip_loss + vf_coef * vf_loss).backward()
    n = logp.numel()
    assert np.allclose(logp.grad.cpu().numpy(), c["dlogp"], rtol=1e-4, atol=1e-6 / n)
    assert np.allclose(value.grad.cpu().numpy(), c["dvalue"], rtol=1e-4, atol=1e-6 / n)
    if norm_adv:
        a = g["in"]["adv"].astype(np.float64)
        assert np.isclose(stats[0].item(), a.mean(), rtol=1e-5, atol=1e-6)
        assert np.isclose(stats[1].item(), a.std(ddof=1), rtol=1e-5)


@pytest.mark.parametrize("m", [1000, 64, 40000])
def test_minibatch_index_gather_and_large_path(m):
    """mb_index gathers the per-sample constants inside the kernel; m > 16384 takes the
    two-launch path.  Checked against the numpy oracle."""
    import tianshou_fork_b200 as ts
    rng = np.random.default_rng(m)
    N = 3 * m
    full = {k: rng.standard_normal(N).astype(np.float32) for k in ("adv", "returns", "v_s")}
    full["logp_old"] = rng.normal(-3, 0.3, N).astype(np.float32)
    mb = rng.permutation(N)[:m]
    logp = (full["logp_old"][mb] + rng.normal(0, 0.2, m)).astype(np.float32)
    value = (full["v_s"][mb] + rng.normal(0, 0.3, m)).astype(np.float32)
    d = {k: torch.from_numpy(v).cuda() for k, v in full.items()}
    lp = torch.from_numpy(logp).cuda().requires_grad_(True)
    val = torch.from_numpy(value).cuda().requires_grad_(True)
    cl, vl, _ = ts.ppo_loss(lp, val, d["logp_old"], d["adv"], d["returns"], d["v_s"],
                            torch.from_numpy(mb).cuda(), 0.2, 3.0, True, True)
    ref = orc.ppo_loss(logp, full["logp_old"][mb], full["adv"][mb], full["returns"][mb],
                       full["v_s"][mb], value, 0.2, 3.0, True, True)
    assert np.isclose(cl.item(), ref[0], rtol=1e-5, atol=1e-7)
    assert np.isclose(vl.item(), ref[1], rtol=1e-5)
    (cl + vl).backward()
    assert np.allclose(lp.grad.cpu().numpy(), ref[2], rtol=1e-4, atol=1e-7 / m)
    assert np.allclose(val.grad.cpu().numpy(), ref[3], rtol=1e-4, atol=1e-7 / m)


class _Actor(torch.nn.Module):
    def __init__(self, obs_dim, n_act):
        super().__init__()
        self.net = torch.nn.Sequential(torch.nn.Linear(obs_dim, 64), torch.nn.Tanh(),
                                       torch.nn.Linear(64, n_act))

    def forward(self, obs, state=None, info=None):
        return self.net(obs), state


def test_cartpole_shaped_ppo_update_end_to_end():
    """20 sub-buffers x 100 steps, obs 4, 2 actions (test/discrete/test_ppo.py:29-33 shapes):
    device rollout -> update(0, buffer) -> losses; compared with an eager-torch evaluation of
    the reference's learn() arithmetic driven by the same minibatch permutations."""
    import tianshou_fork_b200 as ts
    torch.manual_seed(0)
    rng = np.random.default_rng(0)
    E, T = 20, 100
    buf = ts.VectorReplayBuffer(E * T, E)
    for t in range(T):
        term = rng.random(E) < 0.03
        buf.add(ts.Batch(obs=rng.standard_normal((E, 4)).astype(np.float32),
                         act=rng.integers(0, 2, E), rew=np.ones(E), terminated=term,
                         truncated=np.zeros(E, bool),
                         obs_next=rng.standard_normal((E, 4)).astype(np.float32), info={}))

    def make():
        torch.manual_seed(1)
        actor = _Actor(4, 2).cuda()
        critic = torch.nn.Sequential(torch.nn.Linear(4, 64), torch.nn.Tanh(),
                                     torch.nn.Linear(64, 1)).cuda()
        optim = torch.optim.Adam(list(actor.parameters()) + list(critic.parameters()), lr=3e-4)
        return actor, critic, optim

    actor, critic, optim = make()
    dist_fn = lambda logits: torch.distributions.Categorical(logits=logits)
    policy = ts.PPOPolicy(actor, critic, optim, dist_fn, eps_clip=0.2, value_clip=True,
                          advantage_normalization=True, vf_coef=0.5, ent_coef=0.0,
                          max_grad_norm=0.5, gae_lambda=0.95, discount_factor=0.99)
    np.random.seed(7)
    result = policy.update(0, buf, batch_size=64, repeat=2)
    n_mb = 2 * len(ts.policy.ppo._split_bounds(E * T, 64, True))
    assert len(result["loss"]) == n_mb and all(np.isfinite(result["loss"]))

    # ---- eager re-evaluation of ppo.py:99-162 with the same seeds --------------------------
    actor2, critic2, optim2 = make()
    batch, indices = buf.sample(0)
    with torch.no_grad():
        v_s = critic2(batch.obs).flatten()
        v_n = critic2(batch.obs_next).flatten()
    idx = indices.cpu().numpy()
    ret, adv = orc.compute_episodic_return(
        batch.rew.cpu().numpy(), batch.terminated.cpu().numpy(), batch.truncated.cpu().numpy(),
        buf.terminated.cpu().numpy()[idx], idx, buf.unfinished_index(),
        v_s_=v_n.cpu().numpy(), v_s=v_s.cpu().numpy(), gamma=0.99, gae_lambda=0.95)
    ret_t = torch.from_numpy(ret).cuda().float()
    adv_t = torch.from_numpy(adv).cuda().float()
    with torch.no_grad():
        logp_old = dist_fn(actor2(batch.obs)[0]).log_prob(batch.act)
    np.random.seed(7)
    ref_losses = []
    params = list(actor2.parameters()) + list(critic2.parameters())
    for _ in range(2):
        perm = np.random.permutation(E * T)
        for lo, hi in ts.policy.ppo._split_bounds(E * T, 64, True):
            mb = torch.from_numpy(perm[lo:hi]).cuda()
            dist = dist_fn(actor2(batch.obs[mb])[0])
            a = adv_t[mb]
            a = (a - a.mean()) / (a.std() + 1e-8)
            ratio = (dist.log_prob(batch.act[mb]) - logp_old[mb]).exp().float()
            surr1, surr2 = ratio * a, ratio.clamp(0.8, 1.2) * a
            clip_loss = -torch.min(surr1, surr2).mean()
            value = critic2(batch.obs[mb]).flatten()
            v_clip = v_s[mb] + (value - v_s[mb]).clamp(-0.2, 0.2)
            vf_loss = torch.max((ret_t[mb] - value).pow(2), (ret_t[mb] - v_clip).pow(2)).mean()
            loss = clip_loss + 0.5 * vf_loss - 0.0 * dist.entropy().mean()
            optim2.zero_grad()
            loss.backward()
            torch.nn.utils.clip_grad_norm_(params, max_norm=0.5)
            optim2.step()
            ref_losses.append(loss.item())
    assert np.allclose(result["loss"], ref_losses, rtol=2e-3, atol=1e-4)
    assert np.allclose(result["loss"][:5], ref_losses[:5], rtol=1e-4, atol=1e-5)
    for p, q in zip(policy._params, params):
        assert torch.allclose(p, q, rtol=1e-3, atol=1e-4)


# ---- PPOPolicy.update against the reference's own update (tests/golden/ppo_update.npz) -----------
def _load_nets(case):
    from helpers import ppo_update_nets
    Actor, Critic = ppo_update_nets(torch)
    actor, critic = Actor(), Critic()
    for net, tag in ((actor, "actor"), (critic, "critic")):
        sd = {k[len(f"init/{tag}/"):]: torch.from_numpy(v) for k, v in case.items()
              if k.startswith(f"init/{tag}/")}
        net.load_state_dict(sd)
    return actor.cuda(), critic.cuda()


@pytest.mark.parametrize("graph", [False, True], ids=["eager", "cuda_graph"])
@pytest.mark.parametrize("name", ["clip_norm", "rewnorm_recompute_dual"])
def test_ppo_update_vs_reference_update(name, graph):
    """The reference's ``PPOPolicy.update`` (run on CPU by make_golden.gen_ppo_update) against
    this package's on the same rollout, initial weights and numpy seed: returns / advantages /
    logp_old of process_fn, every minibatch's logged losses, the running return statistics
    and the parameters after the update(s).  Tolerances: f32 networks on two devices."""
    import tianshou_fork_b200 as ts
    from helpers import PPO_UPDATE_SPECS, fill
    case = load_golden("ppo_update")[name]
    _, kw, batch_size, repeat, n_updates = next(s for s in PPO_UPDATE_SPECS if s[0] == name)
    buf = ts.VectorReplayBuffer(20 * 100, 20)
    fill(buf, ts.Batch, 700, 100, 20, 0.03, 0.01)
    actor, critic = _load_nets(case)
    optim = torch.optim.Adam(list(actor.parameters()) + list(critic.parameters()), lr=3e-4)
    policy = ts.PPOPolicy(actor, critic, optim,
                          lambda logits: torch.distributions.Categorical(logits=logits),
                          action_scaling=False, cuda_graph=graph, **kw)
    seen = []
    inner = policy.process_fn

    def spy(batch, buffer, indices):
        batch = inner(batch, buffer, indices)
        seen.append({k: batch[k].detach().cpu().numpy().copy()
                     for k in ("returns", "adv", "v_s", "logp_old")})
        return batch

    policy.process_fn = spy
    np.random.seed(7)
    for u in range(n_updates):
        res = policy.update(0, buf, batch_size=batch_size, repeat=repeat)
        tol = 1e-4 if u == 0 else 2e-3  # the second update starts from updated weights
        for k in ("returns", "adv", "v_s", "logp_old"):
            assert np.allclose(seen[u][k], case[f"u{u}/{k}"], rtol=tol, atol=tol), (u, k)
        for k in ("loss", "loss/clip", "loss/vf", "loss/ent"):
            got, want = np.asarray(res[k]), case[f"u{u}/{k}"]
            assert got.shape == want.shape
            assert np.allclose(got, want, rtol=5e-3, atol=5e-4), (u, k)
            if u == 0:
                assert np.allclose(got[:4], want[:4], rtol=1e-4, atol=1e-5), (u, k)
        mean, var, count = case[f"u{u}/ret_rms"]
        assert policy.ret_rms.count == int(count)
        assert np.isclose(policy.ret_rms.mean.item(), mean, rtol=1e-3, atol=1e-4)
        assert np.isclose(policy.ret_rms.var.item(), var, rtol=1e-3)
    for net, tag in ((actor, "actor"), (critic, "critic")):
        for k, v in net.state_dict().items():
            assert np.allclose(v.cpu().numpy(), case[f"final/{tag}/{k}"], rtol=2e-3, atol=2e-4), k


def test_graphed_learn_equals_eager_learn():
    """cuda_graph=True replays the eager minibatch step kernel for kernel: same losses and
    parameters (MuJoCo-like continuous actions, merged tail, two learn calls, lr change)."""
    import tianshou_fork_b200 as ts
    E, T, D, A = 8, 50, 11, 3
    rng = np.random.default_rng(3)
    buf = ts.VectorReplayBuffer(E * T, E)
    for t in range(T):
        buf.add(ts.Batch(obs=rng.standard_normal((E, D)).astype(np.float32),
                         act=rng.standard_normal((E, A)).astype(np.float32),
                         rew=rng.standard_normal(E), terminated=rng.random(E) < 0.05,
                         truncated=np.zeros(E, bool),
                         obs_next=rng.standard_normal((E, D)).astype(np.float32), info={}))

    class Actor(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.mu = torch.nn.Sequential(torch.nn.Linear(D, 32), torch.nn.Tanh(),
                                          torch.nn.Linear(32, A))
            self.sigma = torch.nn.Parameter(torch.zeros(A))

        def forward(self, obs, state=None, info=None):
            mu = self.mu(obs)
            return (mu, self.sigma.exp().expand_as(mu)), state

    def dist_fn(mu, sigma):
        return torch.distributions.Independent(torch.distributions.Normal(mu, sigma), 1)

    out = []
    for graph in (False, True):
        torch.manual_seed(5)
        actor = Actor().cuda()
        critic = torch.nn.Sequential(torch.nn.Linear(D, 32), torch.nn.Tanh(),
                                     torch.nn.Linear(32, 1)).cuda()
        optim = torch.optim.Adam(list(actor.parameters()) + list(critic.parameters()), lr=1e-3)
        policy = ts.PPOPolicy(actor, critic, optim, dist_fn, value_clip=True, max_grad_norm=0.5,
                              cuda_graph=graph)
        np.random.seed(11)
        r1 = policy.update(0, buf, batch_size=96, repeat=3)  # 400 = 3 x 96 + merged 112
        for g in optim.param_groups:
            g["lr"] = 5e-4  # what an lr scheduler does between updates
        r2 = policy.update(0, buf, batch_size=96, repeat=2)
        out.append((r1, r2, [p.detach().clone() for p in policy._params]))
    (e1, e2, ep), (g1, g2, gp) = out
    assert len(g1["loss"]) == 12 and len(g2["loss"]) == 8
    for a, b in ((e1, g1), (e2, g2)):
        for k in a:
            assert np.allclose(a[k], b[k], rtol=1e-5, atol=1e-6), k
    for p, q in zip(ep, gp):
        assert torch.allclose(p, q, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("name", ["pg", "a2c", "a2c_cuda_graph"])
def test_pg_a2c_update_vs_reference_update(name):
    """``PGPolicy.update`` / ``A2CPolicy.update`` of the reference (make_golden.gen_pg_a2c_update,
    return normalisation on, two updates) against this package's device-resident versions."""
    import tianshou_fork_b200 as ts
    from helpers import fill
    graph = name.endswith("_cuda_graph")
    name = name.split("_")[0]
    case = load_golden("pg_a2c_update")[name]
    buf = ts.VectorReplayBuffer(20 * 100, 20)
    fill(buf, ts.Batch, 700, 100, 20, 0.03, 0.01)
    actor, critic = _load_nets(case)
    dist_fn = lambda logits: torch.distributions.Categorical(logits=logits)  # noqa: E731
    if name == "pg":
        optim = torch.optim.Adam(actor.parameters(), lr=1e-3)
        policy = ts.PGPolicy(actor, optim, dist_fn, discount_factor=0.95,
                             reward_normalization=True, action_scaling=False)
        keys = ("returns",)
    else:
        optim = torch.optim.Adam(list(actor.parameters()) + list(critic.parameters()), lr=1e-3)
        policy = ts.A2CPolicy(actor, critic, optim, dist_fn, vf_coef=0.5, ent_coef=0.01,
                              max_grad_norm=0.5, gae_lambda=0.9, discount_factor=0.98,
                              reward_normalization=True, action_scaling=False,
                              cuda_graph=graph)
        keys = ("returns", "adv", "v_s")
    seen = []
    inner = policy.process_fn

    def spy(batch, buffer, indices):
        batch = inner(batch, buffer, indices)
        seen.append({k: batch[k].detach().cpu().numpy().copy() for k in keys})
        return batch

    policy.process_fn = spy
    np.random.seed(9)
    for u in range(2):
        res = policy.update(0, buf, batch_size=100, repeat=2)
        tol = 1e-5 if (u == 0 and name == "pg") else (1e-4 if u == 0 else 3e-3)
        for k in keys:
            assert np.allclose(seen[u][k], case[f"u{u}/{k}"], rtol=tol, atol=tol), (u, k)
        for k in res:
            got, want = np.asarray(res[k]), case[f"u{u}/{k}"]
            assert got.shape == want.shape
            assert np.allclose(got, want, rtol=5e-3, atol=5e-4), (u, k)
        mean, var, count = case[f"u{u}/ret_rms"]
        assert policy.ret_rms.count == int(count)
        assert np.isclose(policy.ret_rms.mean.item(), mean, rtol=1e-3, atol=1e-4)
        assert np.isclose(policy.ret_rms.var.item(), var, rtol=1e-3)
    for net, tag in ((actor, "actor"), (critic, "critic")):
        for k, v in net.state_dict().items():
            assert np.allclose(v.cpu().numpy(), case[f"final/{tag}/{k}"], rtol=3e-3, atol=3e-4), k


@pytest.mark.parametrize("graph", [False, True], ids=["eager", "cuda_graph"])
def test_ppo_learns_the_point_task(graph):
    """End to end like test/discrete/test_ppo.py (collect -> update -> reset, a reward
    threshold at the end): rollouts through add(), GAE and the fused loss on the device."""
    import importlib.util
    import os
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                        "tests", "ppo_point_task.py")
    spec = importlib.util.spec_from_file_location("ppo_point", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    hist = mod.run(iterations=25, graph=graph, seed=1)
    first, last = np.mean(hist[:3]), np.mean(hist[-3:])
    assert last > first + 0.25, (first, last)  # a random policy drifts around -0.95
    assert last > -0.6, hist[-3:]


@pytest.mark.parametrize("graph", [False, True], ids=["eager", "cuda_graph"])
def test_ppo_ant_shaped_update_vs_reference_update(graph):
    """BASELINE.json configs[2] at full size: the reference's ``PPOPolicy.update`` (run on CPU by
    make_golden.gen_ppo_ant) on an Ant-shaped rollout -- obs 27, act 8, 64 envs x 2048 steps =
    131 072 transitions, Gaussian actor, minibatch 1024, one repeat (128 minibatches) -- against
    this package's, same initial weights and numpy seed.  Tolerances: f32 networks on two
    devices; GAE over 2048-step streams in f64 on both sides."""
    import tianshou_fork_b200 as ts
    import helpers as H
    case = load_golden("ppo_ant")["ant"]
    A = H.ANT
    obs, act, rew, term, trunc = H.ant_rollout()
    buf = ts.VectorReplayBuffer(A["E"] * A["T"], A["E"])
    for t in range(A["T"]):
        buf.add(ts.Batch(obs=obs[t], act=act[t], rew=rew[t], terminated=term[t],
                         truncated=trunc[t], obs_next=obs[t + 1], info={}))
    Actor, Critic = H.ant_nets(torch)
    actor, critic = Actor(), Critic()
    for net, tag in ((actor, "actor"), (critic, "critic")):
        net.load_state_dict({k[len(f"init/{tag}/"):]: torch.from_numpy(v)
                             for k, v in case.items() if k.startswith(f"init/{tag}/")})
    actor, critic = actor.cuda(), critic.cuda()
    optim = torch.optim.Adam(list(actor.parameters()) + list(critic.parameters()), lr=3e-4)
    policy = ts.PPOPolicy(actor, critic, optim, H.ant_dist(torch), action_scaling=False,
                          cuda_graph=graph, **H.ANT_KW)
    seen = {}
    inner = policy.process_fn

    def spy(batch, buffer, indices):
        assert buffer.is_canonical(indices) == (A["E"], A["T"])  # the rows kernel's layout
        batch = inner(batch, buffer, indices)
        for k in ("returns", "adv", "v_s", "logp_old"):
            seen[k] = batch[k].detach().double().cpu().numpy().copy()
        return batch

    policy.process_fn = spy
    np.random.seed(13)
    res = policy.update(0, buf, batch_size=A["batch_size"], repeat=A["repeat"])
    for k in ("returns", "adv", "v_s", "logp_old"):
        v = seen[k]
        assert v.shape == (A["E"] * A["T"],)
        assert np.allclose(v[::37], case[f"{k}_sample"], rtol=1e-4, atol=1e-4), k
        stats = np.array([v.sum(), np.abs(v).sum(), (v * v).sum()])
        # sums of 131 072 terms: tolerance tied to the sum of magnitudes, not to the (cancelling) sum
        assert np.allclose(stats, case[f"{k}_stats"], rtol=1e-4,
                           atol=1e-5 * case[f"{k}_stats"][1]), k
    n_mb = A["E"] * A["T"] // A["batch_size"]
    for k in ("loss", "loss/clip", "loss/vf", "loss/ent"):
        got, want = np.asarray(res[k]), case[k]
        assert got.shape == want.shape == (n_mb * A["repeat"],)
        assert np.allclose(got[:4], want[:4], rtol=1e-4, atol=1e-5), k
        assert np.allclose(got, want, rtol=5e-3, atol=5e-4), k
    for net, tag in ((actor, "actor"), (critic, "critic")):
        for k, v in net.state_dict().items():
            assert np.allclose(v.cpu().numpy(), case[f"final/{tag}/{k}"], rtol=3e-3, atol=3e-4), k
