"""DQNPolicy.update on a PrioritizedVectorReplayBuffer against the reference's own update
(tests/golden/dqn_update.npz, written by make_golden.gen_dqn_update): the whole C4 pipeline
sample (PER) -> n-step return -> learn -> update_weight, several updates in a row."""
import numpy as np
import pytest
import torch

from conftest import load_golden

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("graph", [False, True], ids=["eager", "cuda_graph"])
@pytest.mark.parametrize("name", ["double_n3", "nature_n1_huber"])
def test_dqn_per_update_vs_reference_update(name, graph):
    import tianshou_fork_b200 as ts
    from helpers import DQN_UPDATE_SPECS, fill, ppo_update_nets
    case = load_golden("dqn_update")[name]
    _, kw, per_kw, bsz, n_updates = next(s for s in DQN_UPDATE_SPECS if s[0] == name)
    buf = ts.PrioritizedVectorReplayBuffer(20 * 100, 20, **per_kw)
    fill(buf, ts.Batch, 800, 100, 20, 0.03, 0.01)
    Actor, _ = ppo_update_nets(torch)
    net = Actor()
    net.load_state_dict({k[len("init/actor/"):]: torch.from_numpy(v) for k, v in case.items()
                         if k.startswith("init/actor/")})
    net = net.cuda()
    optim = torch.optim.Adam(net.parameters(), lr=1e-3)
    policy = ts.DQNPolicy(net, optim, cuda_graph=graph, **kw)
    seen = []
    inner = policy.process_fn

    def spy(batch, buffer, indices):
        rec = {"indices": indices.cpu().numpy().copy(),
               "weight_in": batch.weight.cpu().numpy().copy()}
        batch = inner(batch, buffer, indices)
        rec["returns"] = batch.returns.detach().cpu().numpy().copy()
        seen.append(rec)
        return batch

    if not graph:  # the spy reads device values on the host, which a captured body cannot
        policy.process_fn = spy
    np.random.seed(13)
    for u in range(n_updates):
        res = policy.update(bsz, buf)
        if not graph:
            # identical host uniforms + (nearly) identical leaves => identical sampled indices
            assert np.array_equal(seen[u]["indices"], case[f"u{u}/indices"]), u
            assert np.allclose(seen[u]["weight_in"], case[f"u{u}/weight_in"], rtol=1e-4,
                               atol=1e-6)
            assert seen[u]["returns"].shape == case[f"u{u}/returns"].shape
            assert np.allclose(seen[u]["returns"], case[f"u{u}/returns"], rtol=1e-3, atol=1e-4)
        assert np.isclose(res["loss"], case[f"u{u}/loss"], rtol=1e-3, atol=1e-5)
        leaves = buf.weight[torch.from_numpy(case[f"u{u}/indices"]).cuda()]
        leaves = leaves.cpu().numpy() if isinstance(leaves, torch.Tensor) else np.asarray(leaves)
        assert np.allclose(leaves, case[f"u{u}/leaves"], rtol=1e-3, atol=1e-4)
        assert np.allclose([buf._max_prio, buf._min_prio], case[f"u{u}/prio"], rtol=1e-3,
                           atol=1e-4)
    for k, v in net.state_dict().items():
        assert np.allclose(v.cpu().numpy(), case[f"final/actor/{k}"], rtol=3e-3, atol=3e-4), k


def test_graphed_update_sees_transitions_added_between_updates():
    """cuda_graph=True on a uniform VectorReplayBuffer that keeps growing: the replayed graph
    reads the refreshed ring bookkeeping, so eager and graphed runs stay identical."""
    import tianshou_fork_b200 as ts
    from helpers import ppo_update_nets
    Actor, _ = ppo_update_nets(torch)
    out = []
    for graph in (False, True):
        rng = np.random.default_rng(5)
        buf = ts.VectorReplayBuffer(8 * 64, 8)
        torch.manual_seed(4)
        net = Actor().cuda()
        optim = torch.optim.Adam(net.parameters(), lr=1e-3)
        policy = ts.DQNPolicy(net, optim, 0.95, estimation_step=2, target_update_freq=3,
                              cuda_graph=graph)
        np.random.seed(21)
        losses = []
        for t in range(40):
            term = rng.random(8) < 0.1
            buf.add(ts.Batch(obs=rng.standard_normal((8, 3)).astype(np.float32),
                             act=rng.integers(0, 4, 8), rew=rng.standard_normal(8),
                             terminated=term, truncated=np.zeros(8, bool),
                             obs_next=rng.standard_normal((8, 3)).astype(np.float32), info={}))
            if t >= 4:
                losses.append(policy.update(16, buf)["loss"])
        out.append((losses, [p.detach().clone() for p in net.parameters()]))
    (le, pe), (lg, pg) = out
    assert np.allclose(le, lg, rtol=1e-5, atol=1e-6)
    for a, b in zip(pe, pg):
        assert torch.allclose(a, b, rtol=1e-5, atol=1e-6)


class _QNet(torch.nn.Module):  # module level: picklable
    def __init__(self):
        super().__init__()
        self.net = torch.nn.Sequential(torch.nn.Linear(3, 32), torch.nn.Tanh(),
                                       torch.nn.Linear(32, 4))

    def forward(self, obs, state=None, info=None):
        return self.net(obs), state


def test_policy_with_captured_graphs_pickles():
    """A policy that has captured CUDA graphs pickles (the graphs are dropped and re-captured)."""
    import pickle
    import tianshou_fork_b200 as ts
    rng = np.random.default_rng(1)
    buf = ts.VectorReplayBuffer(4 * 32, 4)
    for _ in range(32):
        buf.add(ts.Batch(obs=rng.standard_normal((4, 3)).astype(np.float32),
                         act=rng.integers(0, 4, 4), rew=rng.standard_normal(4),
                         terminated=rng.random(4) < 0.1, truncated=np.zeros(4, bool),
                         obs_next=rng.standard_normal((4, 3)).astype(np.float32), info={}))
    net = _QNet().cuda()
    policy = ts.DQNPolicy(net, torch.optim.Adam(net.parameters(), lr=1e-3), 0.9,
                          estimation_step=2, cuda_graph=True)
    np.random.seed(0)
    a = policy.update(8, buf)["loss"]
    clone = pickle.loads(pickle.dumps(policy))
    assert clone._graphed == {} and np.isfinite(a)
    np.random.seed(5)
    l1 = policy.update(8, buf)["loss"]
    np.random.seed(5)
    l2 = clone.update(8, buf)["loss"]  # re-captures against its own copies of the networks
    assert np.isclose(l1, l2, rtol=1e-5)


@pytest.mark.parametrize("prio", [False, True], ids=["ring", "prioritized_ring"])
def test_graphed_update_on_a_growing_single_ring(prio):
    """cuda_graph=True on a plain (Prioritized)ReplayBuffer that is still filling: a single ring
    takes ``%`` by its CURRENT size (base.py:158-168), which the captured kernels read from the
    device bookkeeping -- the by-value ``sub_size`` launch argument is frozen at capture time and
    must not be used for it.  Eager and graphed runs stay identical while the ring grows, wraps
    and keeps going."""
    import tianshou_fork_b200 as ts
    from helpers import ppo_update_nets
    Actor, _ = ppo_update_nets(torch)
    out = []
    for graph in (False, True):
        rng = np.random.default_rng(6)
        buf = ts.PrioritizedReplayBuffer(48, 0.6, 0.4) if prio else ts.ReplayBuffer(48)
        torch.manual_seed(4)
        net = Actor().cuda()
        optim = torch.optim.Adam(net.parameters(), lr=1e-3)
        policy = ts.DQNPolicy(net, optim, 0.95, estimation_step=3, target_update_freq=3,
                              cuda_graph=graph)
        np.random.seed(22)
        losses, nexts = [], []
        for t in range(70):
            buf.add(ts.Batch(obs=rng.standard_normal(3).astype(np.float32),
                             act=int(rng.integers(0, 4)), rew=float(rng.standard_normal()),
                             terminated=bool(rng.random() < 0.1), truncated=False,
                             obs_next=rng.standard_normal(3).astype(np.float32), info={}))
            if t >= 6:
                losses.append(policy.update(8, buf)["loss"])
                nexts.append(buf.next(np.arange(len(buf))))
        out.append((losses, nexts, [p.detach().clone() for p in net.parameters()]))
    (le, ne, pe), (lg, ng, pg) = out
    assert np.allclose(le, lg, rtol=1e-5, atol=1e-6)
    assert all(np.array_equal(a, b) for a, b in zip(ne, ng))
    for a, b in zip(pe, pg):
        assert torch.allclose(a, b, rtol=1e-5, atol=1e-6)
