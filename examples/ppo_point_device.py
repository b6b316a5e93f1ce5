#!/usr/bin/env python
"""The same on-policy loop as ``ppo_point.py`` with the simulator ON the GPU: 4096 env streams
stepped by a few torch ops, every vectorised step stored by ONE launch
(``ReplayBufferManager.add_device``), GAE by the canonical rows kernel, the
PPO minibatch steps replayed from a CUDA graph.  Nothing but the logged losses crosses PCIe.

    python examples/ppo_point_device.py [--envs 4096] [--steps 32] [--iterations 20]
"""
from __future__ import annotations

import argparse
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from ppo_point import MOVES, Actor, ts  # noqa: E402


def run(iterations: int = 20, envs: int = 4096, steps: int = 32, horizon: int = 16,
        graph: bool = True, seed: int = 0):
    """Returns (mean step reward per iteration, env steps per second over the timed part)."""
    torch.manual_seed(seed)
    np.random.seed(seed)
    dev = torch.device("cuda:0")
    moves = torch.from_numpy(MOVES).to(dev)
    actor = Actor().to(dev)
    critic = torch.nn.Sequential(torch.nn.Linear(2, 64), torch.nn.Tanh(),
                                 torch.nn.Linear(64, 1)).to(dev)
    optim = torch.optim.Adam(list(actor.parameters()) + list(critic.parameters()), lr=3e-3)
    policy = ts.PPOPolicy(actor, critic, optim,
                          lambda logits: torch.distributions.Categorical(logits=logits),
                          discount_factor=0.95, gae_lambda=0.95, vf_coef=0.5, ent_coef=0.01,
                          max_grad_norm=0.5, value_clip=True, max_batchsize=envs * steps,
                          cuda_graph=graph)
    pos = torch.rand(envs, 2, device=dev) * 2 - 1
    age = torch.zeros(envs, dtype=torch.int64, device=dev)
    never = torch.zeros(envs, dtype=torch.bool, device=dev)
    buf = ts.VectorReplayBuffer(envs * steps, envs, rew_dtype=torch.float32)
    history, t0, timed_from = [], None, 2
    for it in range(iterations):
        if it == timed_from:  # the first iterations allocate storage and capture the graph
            torch.cuda.synchronize()
            t0 = time.perf_counter()
        rewards = torch.zeros((), device=dev)
        for _ in range(steps):
            obs = pos
            with torch.no_grad():
                act = policy(ts.Batch(obs=obs, info={})).act
            nxt = (pos + moves[act]).clamp(-2, 2)
            age = age + 1
            rew = -nxt.norm(dim=1)
            trunc = age >= horizon
            buf.add_device(ts.Batch(obs=obs, act=act, rew=rew, terminated=never,
                                    truncated=trunc, obs_next=nxt))
            pos = torch.where(trunc[:, None], torch.rand_like(nxt) * 2 - 1, nxt)
            age = torch.where(trunc, torch.zeros_like(age), age)
            rewards = rewards + rew.mean()
        policy.update(0, buf, batch_size=envs * steps // 4, repeat=2)
        buf.reset()
        history.append(rewards / steps)
    torch.cuda.synchronize()
    rate = (iterations - timed_from) * envs * steps / (time.perf_counter() - t0)
    return [float(h) for h in history], rate


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--envs", type=int, default=4096)
    ap.add_argument("--steps", type=int, default=32)
    ap.add_argument("--iterations", type=int, default=20)
    ap.add_argument("--eager", action="store_true")
    a = ap.parse_args()
    hist, rate = run(a.iterations, a.envs, a.steps, graph=not a.eager)
    for i, r in enumerate(hist):
        print(f"iteration {i:3d}  mean step reward {r:8.4f}")
    print(f"{rate / 1e6:.2f} M env steps per second (collect + add_device + GAE + PPO update)")
