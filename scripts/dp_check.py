#!/usr/bin/env python
"""2+ GPU check of the data-parallel learners (learner row):

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 scripts/dp_check.py

Every rank owns its own env streams (different rollouts), the replicas start from the same
weights, gradients are averaged with one NCCL all-reduce per minibatch.  Checks: (1) after
PPOPolicy.update the parameters are bit-identical on every rank, eager and CUDA-graphed;
(2) the graphed run equals the eager run; (3) with identical rollouts on every rank the
data-parallel update equals the single-process update.  Prints one JSON line on rank 0."""
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tianshou_fork_b200 as ts  # noqa: E402

E, T, D, A = 16, 128, 27, 8


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


def dist_fn(mu, sigma):
    return torch.distributions.Independent(torch.distributions.Normal(mu, sigma), 1)


def rollout(seed):
    rng = np.random.default_rng(seed)
    buf = ts.VectorReplayBuffer(E * T, E)
    for _ in range(T):
        buf.add(ts.Batch(obs=rng.standard_normal((E, D)).astype(np.float32),
                         act=rng.standard_normal((E, A)).astype(np.float32),
                         rew=rng.standard_normal(E), terminated=rng.random(E) < 0.02,
                         truncated=np.zeros(E, bool),
                         obs_next=rng.standard_normal((E, D)).astype(np.float32), info={}))
    return buf


def run(buf, graph, parallel):
    torch.manual_seed(0)
    actor = Actor().cuda()
    critic = torch.nn.Sequential(torch.nn.Linear(D, 64), torch.nn.Tanh(),
                                 torch.nn.Linear(64, 1)).cuda()
    optim = torch.optim.Adam(list(actor.parameters()) + list(critic.parameters()), lr=3e-4)
    pol = ts.PPOPolicy(actor, critic, optim, dist_fn, value_clip=True, max_grad_norm=0.5,
                       cuda_graph=graph)
    if parallel:
        pol.set_data_parallel(True)
    np.random.seed(3)  # same minibatch permutation on every rank
    pol.update(0, buf, batch_size=256, repeat=1)  # capture + first update
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    res = pol.update(0, buf, batch_size=256, repeat=2)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    flat = torch.cat([p.detach().reshape(-1) for p in pol._params])
    return flat, res, dt / len(res["loss"]) * 1e6


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dist.init_process_group("nccl")
    out = {"world": world}
    mine = rollout(100 + rank)
    for graph in (False, True):
        flat, res, us = run(mine, graph, True)
        gathered = [torch.empty_like(flat) for _ in range(world)]
        dist.all_gather(gathered, flat)
        same = all(torch.equal(gathered[0], g) for g in gathered)
        out["graph" if graph else "eager"] = {"replicas_identical": bool(same),
                                              "us_per_minibatch": us,
                                              "loss0": res["loss"][0]}
        assert same, "replicas diverged"
        if graph:
            assert torch.allclose(flat, eager_flat, rtol=1e-5, atol=1e-6), "graph != eager"
            out["graph_equals_eager"] = True
        eager_flat = flat
    # identical rollouts everywhere: the averaged gradient IS the local gradient
    shared = rollout(7)
    dp, _, _ = run(shared, False, True)
    single, _, _ = run(shared, False, False)
    out["dp_equals_single_on_identical_data"] = bool(torch.allclose(dp, single, rtol=1e-5,
                                                                    atol=1e-6))
    assert out["dp_equals_single_on_identical_data"]
    if rank == 0:
        print(json.dumps(out), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
