"""`updates` leg of bench.py at N > 1: the data-parallel learner (learner row).

BASELINE.json configs[2] shape (Ant: obs 27, act 8, 64 env streams x 2048 steps per rank, two
64 x 64 tanh MLPs, Independent Normal, value clip + advantage normalisation + grad-norm clip +
Adam).  Every rank learns on its OWN rollout; `policy.set_data_parallel()` averages the gradients
with one flat NCCL all-reduce per minibatch (policy/base.py:allreduce_mean_grads).  Reported per
minibatch (wall time of `PPOPolicy.learn` / its minibatches, max over ranks), eager and
`cuda_graph=True`, next to the same learner WITHOUT the all-reduce on the same rank, plus the
latency of the all-reduce alone (eager launches and inside a CUDA graph) and a check that the
replicas end bit-identical.
"""
from __future__ import annotations

import sys
import time

import numpy as np

D, A, E, T = 27, 8, 64, 2048


def _nets(torch, dev):
    class Actor(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.mu = torch.nn.Sequential(torch.nn.Linear(D, 64), torch.nn.Tanh(),
                                          torch.nn.Linear(64, 64), torch.nn.Tanh(),
                                          torch.nn.Linear(64, A))
            self.sigma = torch.nn.Parameter(torch.zeros(A))

        def forward(self, obs, state=None, info=None):
            mu = self.mu(obs)
            return (mu, self.sigma.exp().expand_as(mu)), state

    torch.manual_seed(0)  # the same initial weights on every rank
    actor = Actor().to(dev)
    critic = torch.nn.Sequential(torch.nn.Linear(D, 64), torch.nn.Tanh(),
                                 torch.nn.Linear(64, 64), torch.nn.Tanh(),
                                 torch.nn.Linear(64, 1)).to(dev)
    optim = torch.optim.Adam(list(actor.parameters()) + list(critic.parameters()), lr=3e-4)
    return actor, critic, optim


def _data(torch, n, dev, seed):
    g = torch.Generator().manual_seed(seed)
    r = lambda *sh: torch.randn(*sh, generator=g)  # noqa: E731
    v_s = r(n)
    return dict(obs=r(n, D).to(dev), act=r(n, A).to(dev), v_s=v_s.to(dev), adv=r(n).to(dev),
                returns=(v_s + 0.3 * r(n)).to(dev), logp_old=(-11.0 + 0.3 * r(n)).to(dev))


def allreduce_latency(torch, dist, dev, numel, max_over_ranks):
    """us per all-reduce of `numel` f32 (the learner's flat gradient), eager and graphed."""
    flat = torch.zeros(numel, device=dev)
    for _ in range(20):
        dist.all_reduce(flat)
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(200):
        dist.all_reduce(flat)
    e1.record()
    torch.cuda.synchronize()
    eager = max_over_ranks(e0.elapsed_time(e1) * 1e3 / 200)
    graphed = None
    try:
        side = torch.cuda.Stream(dev)
        with torch.cuda.stream(side):
            dist.all_reduce(flat)
        side.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):
            for _ in range(8):
                dist.all_reduce(flat)
                flat.mul_(0.5)  # a dependent kernel between collectives, like the optimiser step
        for _ in range(3):
            g.replay()
        torch.cuda.synchronize()
        dist.barrier()
        e0.record()
        for _ in range(25):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
        graphed = max_over_ranks(e0.elapsed_time(e1) * 1e3 / 200)
    except Exception as e:  # pragma: no cover
        graphed = f"{type(e).__name__}: {e}"
    return {"numel": numel, "eager_us": eager, "in_cuda_graph_us_incl_one_small_kernel": graphed}


def run_dp_updates(torch, dist, dev, rank, world, max_over_ranks):
    try:
        import tianshou_fork_b200 as ts
        dist_fn = lambda mu, sigma: torch.distributions.Independent(  # noqa: E731
            torch.distributions.Normal(mu, sigma), 1)
        out = {"unit": "us per minibatch (wall time of PPOPolicy.learn / minibatches, max over "
                       "ranks)", "world": world,
               "shape": f"{E} env streams x {T} steps per rank, obs {D}, act {A}"}
        n = E * T
        for bs in (1024, 64):
            count = n // bs if bs == 1024 else 256
            rows = {}
            for mode in ("eager", "cuda_graph"):
                for dp in (True, False):
                    print(f"[bench_dp r{rank}] mb {bs} {mode} dp={dp}", file=sys.stderr, flush=True)
                    actor, critic, optim = _nets(torch, dev)
                    pol = ts.PPOPolicy(actor, critic, optim, dist_fn, eps_clip=0.2,
                                       value_clip=True, vf_coef=0.25, ent_coef=0.0,
                                       max_grad_norm=0.5, cuda_graph=(mode == "cuda_graph"))
                    if dp:
                        pol.set_data_parallel(True)
                    batch = ts.Batch(**_data(torch, bs * count, dev, seed=1 + rank))
                    np.random.seed(5)  # the same permutation on every rank
                    pol.learn(batch, bs, 1)  # warm-up / capture
                    best = 1e30
                    for _ in range(2):
                        torch.cuda.synchronize()
                        dist.barrier()
                        t0 = time.perf_counter()
                        pol.learn(batch, bs, 1)
                        torch.cuda.synchronize()
                        best = min(best, time.perf_counter() - t0)
                    rows[f"{mode}_{'dp' if dp else 'single'}"] = max_over_ranks(best) * 1e6 / count
                    if dp:
                        flat = torch.cat([p.detach().reshape(-1) for p in pol._params])
                        ref = flat.clone()
                        dist.broadcast(ref, 0)
                        same = torch.tensor([float(torch.equal(flat, ref))], device=dev)
                        dist.all_reduce(same, op=dist.ReduceOp.MIN)
                        rows[f"{mode}_replicas_identical"] = bool(same.item() == 1.0)
                        numel = flat.numel()
            rows["graph_dp_over_single"] = rows["cuda_graph_dp"] / rows["cuda_graph_single"]
            out[f"c3_ppo_learn_mb{bs}"] = rows
        out["allreduce"] = allreduce_latency(torch, dist, dev, numel, max_over_ranks)
        return out
    except Exception:
        raise  # every rank is inside a collective sequence: a rank that drops out must fail the run
