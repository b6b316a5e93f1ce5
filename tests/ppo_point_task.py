"""Test infrastructure for tests/test_gpu_ppo.py::test_ppo_learns_the_point_task (the stand-in
for BASELINE.json configs[0], test/discrete/test_ppo.py, which needs gymnasium): an on-policy
loop the way the reference's onpolicy trainer drives it (collect n_step transitions per env ->
policy.update(0, buffer, ...) -> buffer.reset()).

The environment is a vectorised NumPy toy (a point on a plane, five discrete moves, reward
-distance to the origin, fixed-length episodes) so that the example needs no simulator.  The
rollout goes into a ``VectorReplayBuffer`` with ``add`` exactly as ``Collector.collect`` does
(collector.py:324); GAE, the minibatch gather and the PPO loss run in the CUDA kernels.
"""
from __future__ import annotations

import argparse
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tianshou_fork_b200 as ts  # noqa: E402

MOVES = np.array([[0, 0], [1, 0], [-1, 0], [0, 1], [0, -1]], dtype=np.float32) * 0.25


class PointEnvs:
    """E independent points; an episode lasts ``horizon`` steps and then restarts."""

    def __init__(self, n: int, horizon: int, seed: int) -> None:
        self.n, self.horizon = n, horizon
        self.rng = np.random.default_rng(seed)
        self.pos = self.rng.uniform(-1, 1, (n, 2)).astype(np.float32)
        self.t = np.zeros(n, dtype=np.int64)

    def step(self, act: np.ndarray):
        self.pos = np.clip(self.pos + MOVES[act], -2, 2)
        self.t += 1
        rew = -np.linalg.norm(self.pos, axis=1)
        truncated = self.t >= self.horizon
        obs_next = self.pos.copy()
        done = np.flatnonzero(truncated)
        self.pos[done] = self.rng.uniform(-1, 1, (done.size, 2)).astype(np.float32)
        self.t[done] = 0
        return obs_next, rew, np.zeros(self.n, bool), truncated


class Actor(torch.nn.Module):
    def __init__(self):
        super().__init__()
        self.net = torch.nn.Sequential(torch.nn.Linear(2, 64), torch.nn.Tanh(),
                                       torch.nn.Linear(64, 5))

    def forward(self, obs, state=None, info=None):
        return self.net(torch.as_tensor(obs, dtype=torch.float32,
                                        device=self.net[0].weight.device)), state


def run(iterations: int = 30, envs: int = 16, steps: int = 64, graph: bool = False,
        seed: int = 0):
    """Returns the mean step reward of every iteration's rollout."""
    torch.manual_seed(seed)
    np.random.seed(seed)
    dev = torch.device("cuda:0")
    actor = Actor().to(dev)
    critic = torch.nn.Sequential(torch.nn.Linear(2, 64), torch.nn.Tanh(),
                                 torch.nn.Linear(64, 1)).to(dev)
    optim = torch.optim.Adam(list(actor.parameters()) + list(critic.parameters()), lr=3e-3)
    policy = ts.PPOPolicy(actor, critic, optim,
                          lambda logits: torch.distributions.Categorical(logits=logits),
                          discount_factor=0.95, gae_lambda=0.95, vf_coef=0.5, ent_coef=0.01,
                          max_grad_norm=0.5, value_clip=True, max_batchsize=envs * steps,
                          cuda_graph=graph)
    env = PointEnvs(envs, horizon=16, seed=seed)
    buf = ts.VectorReplayBuffer(envs * steps, envs)
    history = []
    for _ in range(iterations):
        obs = env.pos.copy()
        total = 0.0
        for _ in range(steps):
            with torch.no_grad():
                act = policy(ts.Batch(obs=torch.from_numpy(obs).to(dev), info={})).act
            act = act.cpu().numpy()
            obs_next, rew, term, trunc = env.step(act)
            buf.add(ts.Batch(obs=obs, act=act, rew=rew, terminated=term, truncated=trunc,
                             obs_next=obs_next, info={}))
            total += float(rew.mean())
            obs = env.pos.copy()
        history.append(total / steps)
        policy.update(0, buf, batch_size=256, repeat=4)
        buf.reset()
    return history


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--graph", action="store_true", help="PPOPolicy(cuda_graph=True)")
    ap.add_argument("--iterations", type=int, default=30)
    a = ap.parse_args()
    hist = run(a.iterations, graph=a.graph)
    for i, r in enumerate(hist):
        print(f"iteration {i:3d}  mean step reward {r:8.4f}")
