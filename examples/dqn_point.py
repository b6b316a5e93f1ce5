#!/usr/bin/env python
"""Off-policy loop on the device-resident stack, the way the reference's offpolicy trainer
drives it (collect a step per env -> ``policy.update(batch_size, buffer)`` a few times):

    python examples/dqn_point.py [--graph]

Same toy task as ``ppo_point.py``; transitions go into a ``PrioritizedVectorReplayBuffer``
(``add`` initialises the priorities, prio.py:50-57), every update is prioritized sampling ->
3-step return with a target network -> Q-learning step -> ``update_weight`` with the TD errors.
"""
from __future__ import annotations

import argparse
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from ppo_point import Actor, PointEnvs, ts  # noqa: E402


def run(steps: int = 600, envs: int = 16, graph: bool = False, seed: int = 0):
    """Returns the mean step reward of consecutive blocks of 50 collect steps."""
    torch.manual_seed(seed)
    np.random.seed(seed)
    dev = torch.device("cuda:0")
    net = Actor().to(dev)
    optim = torch.optim.Adam(net.parameters(), lr=1e-3)
    policy = ts.DQNPolicy(net, optim, discount_factor=0.9, estimation_step=3,
                          target_update_freq=50, cuda_graph=graph)
    env = PointEnvs(envs, horizon=16, seed=seed)
    buf = ts.PrioritizedVectorReplayBuffer(envs * 512, envs, alpha=0.6, beta=0.4)
    history, block = [], []
    for step in range(steps):
        policy.set_eps(max(0.05, 1.0 - step / (0.5 * steps)))
        obs = env.pos.copy()
        with torch.no_grad():
            act = policy(ts.Batch(obs=torch.from_numpy(obs).to(dev), info={})).act
        act = policy.exploration_noise(act.cpu().numpy(), ts.Batch(obs=obs))
        obs_next, rew, term, trunc = env.step(act)
        buf.add(ts.Batch(obs=obs, act=act, rew=rew, terminated=term, truncated=trunc,
                         obs_next=obs_next, info={}))
        block.append(float(rew.mean()))
        if step >= 20:
            for _ in range(2):
                policy.update(64, buf)
        if len(block) == 50:
            history.append(float(np.mean(block)))
            block = []
    return history


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--graph", action="store_true", help="DQNPolicy(cuda_graph=True)")
    ap.add_argument("--steps", type=int, default=600)
    a = ap.parse_args()
    for i, r in enumerate(run(a.steps, graph=a.graph)):
        print(f"steps {50 * i:4d}-{50 * i + 49:4d}  mean step reward {r:8.4f}")
