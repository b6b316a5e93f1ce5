"""Synchronous advantage actor-critic on device-resident rollouts -- drop-in for
``tianshou.policy.A2CPolicy`` on the return path (tianshou/policy/modelfree/a2c.py:48-175).

``_compute_returns`` (a2c.py:83-117) never leaves the device: no ``.cpu().numpy()`` of the
critic outputs (a2c.py:91-93) and no ``to_torch_as`` copies back (a2c.py:115-116); GAE runs in
``csrc/gae.cu`` and the running return statistics are device scalars.
"""
from __future__ import annotations

from typing import Any, Callable, Dict, List, Optional

import torch
from torch import nn

from ..data.batch import Batch
from .pg import PGPolicy, _minibatches, _split_bounds


class A2CPolicy(PGPolicy):
    """Constructor arguments follow the reference (a2c.py:48-75): ``critic(obs) -> value``."""

    def __init__(
        self,
        actor: nn.Module,
        critic: nn.Module,
        optim: torch.optim.Optimizer,
        dist_fn: Callable[..., torch.distributions.Distribution],
        vf_coef: float = 0.5,
        ent_coef: float = 0.01,
        max_grad_norm: Optional[float] = None,
        gae_lambda: float = 0.95,
        max_batchsize: int = 256,
        **kwargs: Any,
    ) -> None:
        kwargs.setdefault("action_scaling", False)
        kwargs.setdefault("action_bound_method", None)
        super().__init__(actor, optim, dist_fn, **kwargs)
        self.critic = critic
        assert 0.0 <= gae_lambda <= 1.0, "GAE lambda should be in [0, 1]."
        self._lambda = gae_lambda
        self._weight_vf = vf_coef
        self._weight_ent = ent_coef
        self._grad_norm = max_grad_norm
        self._batch = max_batchsize
        seen = {id(p) for p in actor.parameters()}
        self._params = list(actor.parameters()) + [
            p for p in critic.parameters() if id(p) not in seen]  # ActorCritic (a2c.py:75)

    # ---- returns (a2c.py:83-117) ------------------------------------------------------------------
    def _values(self, obs: Any) -> torch.Tensor:
        n = len(obs)
        # Batch.split(self._batch, shuffle=False, merge_last=True)  (a2c.py:88)
        out = [self.critic(obs[lo:hi]) for lo, hi in _split_bounds(n, self._batch, True)]
        return torch.cat(out, dim=0).flatten()

    def _compute_returns(self, batch: Batch, buffer: Any, indices: Any) -> Batch:
        with torch.no_grad():
            batch.v_s = self._values(batch.obs)  # old value
            v_s_ = self._values(batch.obs_next)
        v_s = batch.v_s
        if self._rew_norm:  # unnormalize v_s & v_s_  (a2c.py:98-100)
            scale = torch.sqrt(self.ret_rms.var + self._eps).to(v_s.dtype)
            v_s, v_s_ = v_s * scale, v_s_ * scale
        unnormalized_returns, advantages = self.compute_episodic_return(
            batch, buffer, indices, v_s_, v_s, gamma=self._gamma, gae_lambda=self._lambda)
        if self._rew_norm:  # a2c.py:110-112
            batch.returns = unnormalized_returns / torch.sqrt(
                self.ret_rms.var + self._eps).to(unnormalized_returns.dtype)
            self.ret_rms.update(unnormalized_returns)
        else:
            batch.returns = unnormalized_returns
        batch.returns = batch.returns.to(batch.v_s.dtype)
        batch.adv = advantages.to(batch.v_s.dtype)
        return batch

    def process_fn(self, batch: Batch, buffer: Any, indices: Any) -> Batch:
        batch = self._compute_returns(batch, buffer, indices)
        batch.act = self._act_tensor(batch)
        return batch

    # ---- learn (a2c.py:119-161) -------------------------------------------------------------------
    def learn(self, batch: Batch, batch_size: int, repeat: int, **kwargs: Any
              ) -> Dict[str, List[float]]:
        n = len(batch)
        logged: List[torch.Tensor] = []
        for _ in range(repeat):
            for mb in _minibatches(n, batch_size, self.device):
                obs = batch.obs[mb]
                dist = self._dist(obs)[2]
                adv, ret = batch.adv[mb], batch.returns[mb]
                log_prob = dist.log_prob(batch.act[mb]).reshape(len(adv), -1).transpose(0, 1)
                actor_loss = -(log_prob * adv).mean()
                value = self.critic(obs).flatten()
                vf_loss = nn.functional.mse_loss(ret, value)
                ent_loss = dist.entropy().mean()
                loss = actor_loss + self._weight_vf * vf_loss - self._weight_ent * ent_loss
                self.optim.zero_grad()
                loss.backward()
                self._sync_grads(self._params)
                if self._grad_norm:
                    nn.utils.clip_grad_norm_(self._params, max_norm=self._grad_norm)
                self.optim.step()
                logged.append(torch.stack([loss.detach(), actor_loss.detach(),
                                           vf_loss.detach(), ent_loss.detach()]))
        table = torch.stack(logged).cpu().numpy() if logged else torch.zeros(0, 4).numpy()
        return {
            "loss": table[:, 0].tolist(),
            "loss/actor": table[:, 1].tolist(),
            "loss/vf": table[:, 2].tolist(),
            "loss/ent": table[:, 3].tolist(),
        }
