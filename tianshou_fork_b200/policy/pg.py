"""REINFORCE on device-resident rollouts -- drop-in for ``tianshou.policy.PGPolicy`` on the
return path (tianshou/policy/modelfree/pg.py:45-215): ``process_fn`` computes the Monte-Carlo
returns with the GAE kernel (lambda = 1, ``v_s_ = ret_rms.mean``) and normalises them with
running statistics that stay on the device (tianshou/utils/statistics.py:69-114).
"""
from __future__ import annotations

from typing import Any, Callable, Dict, Iterator, List, Optional, Tuple

import numpy as np
import torch

from ..data.batch import Batch
from .base import BasePolicy


class RunningMeanStd:
    """Running mean / variance of a stream (tianshou/utils/statistics.py:69-114) kept in
    device scalars so that return normalisation (a2c.py:98-112, pg.py:113-117) does not sync."""

    def __init__(self, device: torch.device, mean: float = 0.0, std: float = 1.0,
                 clip_max: Optional[float] = 10.0,
                 epsilon: float = np.finfo(np.float32).eps.item()):
        self.mean = torch.full((), float(mean), dtype=torch.float64, device=device)
        self.var = torch.full((), float(std), dtype=torch.float64, device=device)
        self.clip_max = clip_max
        self.count = 0
        self.eps = epsilon

    def norm(self, x: torch.Tensor) -> torch.Tensor:
        out = (x - self.mean) / torch.sqrt(self.var + self.eps)
        if self.clip_max:
            out = out.clamp(-self.clip_max, self.clip_max)
        return out

    def update(self, x: torch.Tensor) -> None:
        x = x.to(torch.float64)
        batch_var, batch_mean = torch.var_mean(x, dim=0, unbiased=False)
        batch_count = len(x)
        delta = batch_mean - self.mean
        total = self.count + batch_count
        new_mean = self.mean + delta * batch_count / total
        m_a, m_b = self.var * self.count, batch_var * batch_count
        m_2 = m_a + m_b + delta ** 2 * self.count * batch_count / total
        self.mean, self.var, self.count = new_mean, m_2 / total, total


def _split_bounds(n: int, size: int, merge_last: bool) -> List[Tuple[int, int]]:
    """Chunk boundaries of ``Batch.split`` (data/batch.py:896-912)."""
    if size == -1:
        size = n
    assert size >= 1
    merge_last = merge_last and n % size > 0
    out = []
    for idx in range(0, n, size):
        if merge_last and idx + size + size >= n:
            out.append((idx, n))
            break
        out.append((idx, min(idx + size, n)))
    return out


def _minibatches(n: int, size: int, device: torch.device) -> Iterator[torch.Tensor]:
    """Index tensors of ``batch.split(size, merge_last=True)``: the permutation is drawn on
    the host exactly where the reference draws it (data/batch.py:905), so that a seeded run
    visits the reference's minibatches; one H2D copy per epoch."""
    perm = torch.from_numpy(np.random.permutation(n)).to(device)
    for lo, hi in _split_bounds(n, size, merge_last=True):
        yield perm[lo:hi]


class PGPolicy(BasePolicy):
    """Constructor arguments follow the reference (pg.py:45-85): ``model(obs, state=, info=)
    -> (logits, state)``, ``dist_fn(*logits)``."""

    def __init__(
        self,
        model: torch.nn.Module,
        optim: torch.optim.Optimizer,
        dist_fn: Callable[..., torch.distributions.Distribution],
        discount_factor: float = 0.99,
        reward_normalization: bool = False,
        action_scaling: bool = True,
        action_bound_method: Optional[str] = "clip",
        deterministic_eval: bool = False,
        **kwargs: Any,
    ) -> None:
        super().__init__(action_scaling=action_scaling,
                         action_bound_method=action_bound_method, **kwargs)
        self.actor = model
        self.optim = optim
        self.dist_fn = dist_fn
        assert 0.0 <= discount_factor <= 1.0, "discount factor should be in [0, 1]"
        self._gamma = discount_factor
        self._rew_norm = reward_normalization
        self.device = next(model.parameters()).device
        self.ret_rms = RunningMeanStd(self.device)
        self._eps = 1e-8
        self._deterministic_eval = deterministic_eval

    # ---- forward (pg.py:133-172) ------------------------------------------------------------------
    def _dist(self, obs: Any, state: Any = None, info: Any = None):
        logits, hidden = self.actor(obs, state=state, info=info)
        dist = self.dist_fn(*logits) if isinstance(logits, tuple) else self.dist_fn(logits)
        return logits, hidden, dist

    def forward(self, batch: Batch, state: Any = None, **kwargs: Any) -> Batch:
        logits, hidden, dist = self._dist(batch.obs, state, batch.get("info"))
        if self._deterministic_eval and not self.training:
            act = logits[0] if isinstance(logits, tuple) else logits.argmax(-1)  # pg.py:122-131
        else:
            act = dist.sample()
        return Batch(logits=logits, act=act, state=hidden, dist=dist)

    # ---- returns (pg.py:87-120) -------------------------------------------------------------------
    def process_fn(self, batch: Batch, buffer: Any, indices: Any) -> Batch:
        n = len(batch)
        v_s_ = self.ret_rms.mean.expand(n).contiguous()  # np.full(indices.shape, mean)  :108
        unnormalized_returns, _ = self.compute_episodic_return(
            batch, buffer, indices, v_s_=v_s_, gamma=self._gamma, gae_lambda=1.0)
        if self._rew_norm:  # :113-117
            batch.returns = (unnormalized_returns - self.ret_rms.mean) / torch.sqrt(
                self.ret_rms.var + self._eps)
            self.ret_rms.update(unnormalized_returns)
        else:
            batch.returns = unnormalized_returns
        return batch

    def _act_tensor(self, batch: Batch) -> torch.Tensor:
        act = batch.act
        if not isinstance(act, torch.Tensor):
            act = torch.as_tensor(np.asarray(act))
        return act.to(self.device)

    # ---- learn (pg.py:174-192) --------------------------------------------------------------------
    def learn(self, batch: Batch, batch_size: int, repeat: int, **kwargs: Any
              ) -> Dict[str, List[float]]:
        n = len(batch)
        act_all = self._act_tensor(batch)
        ret_all = batch.returns.to(device=self.device, dtype=torch.float32)
        logged: List[torch.Tensor] = []
        for _ in range(repeat):
            for mb in _minibatches(n, batch_size, self.device):
                self.optim.zero_grad()
                dist = self._dist(batch.obs[mb])[2]
                ret = ret_all[mb]
                log_prob = dist.log_prob(act_all[mb]).reshape(len(ret), -1).transpose(0, 1)
                loss = -(log_prob * ret).mean()
                loss.backward()
                self._sync_grads(self.actor.parameters())
                self.optim.step()
                logged.append(loss.detach())
        losses = torch.stack(logged).cpu().tolist() if logged else []  # one sync per learn
        return {"loss": losses}
