"""Deep Q-learning on a device-resident (prioritized) replay buffer -- drop-in for
``tianshou.policy.DQNPolicy`` on the n-step / PER path (tianshou/policy/modelfree/dqn.py:40-215).

The pipeline of one ``update`` never leaves the device: ``buffer.sample`` (PER descent +
importance weights, ``csrc/segtree.cu``) -> ``process_fn`` = ``compute_nstep_return``
(``csrc/nstep.cu``, the target network evaluated on ``buffer[terminal]``) -> ``learn`` ->
``post_process_fn`` = ``buffer.update_weight(indices, td_error)`` (policy/base.py:285-286).
The only host read is the logged ``loss.item()`` the reference API returns (dqn.py:194).
"""
from __future__ import annotations

from copy import deepcopy
from typing import Any, Dict, Optional

import numpy as np
import torch

from ..data.batch import Batch
from .base import BasePolicy


class DQNPolicy(BasePolicy):
    """Constructor arguments follow the reference (dqn.py:40-68):
    ``model(obs, state=, info=) -> (logits, state)``."""

    def __init__(
        self,
        model: torch.nn.Module,
        optim: torch.optim.Optimizer,
        discount_factor: float = 0.99,
        estimation_step: int = 1,
        target_update_freq: int = 0,
        reward_normalization: bool = False,
        is_double: bool = True,
        clip_loss_grad: bool = False,
        **kwargs: Any,
    ) -> None:
        super().__init__(**kwargs)
        self.model = model
        self.optim = optim
        self.eps = 0.0
        assert 0.0 <= discount_factor <= 1.0, "discount factor should be in [0, 1]"
        self._gamma = discount_factor
        assert estimation_step > 0, "estimation_step should be greater than 0"
        self._n_step = estimation_step
        self._target = target_update_freq > 0
        self._freq = target_update_freq
        self._iter = 0
        if self._target:
            self.model_old = deepcopy(self.model)
            self.model_old.eval()
        self._rew_norm = reward_normalization
        self._is_double = is_double
        self._clip_loss_grad = clip_loss_grad
        self.device = next(model.parameters()).device

    def set_eps(self, eps: float) -> None:
        self.eps = eps

    def train(self, mode: bool = True) -> "DQNPolicy":
        """Training mode for everything except the target network (dqn.py:74-78)."""
        self.training = mode
        self.model.train(mode)
        return self

    def sync_weight(self) -> None:
        self.model_old.load_state_dict(self.model.state_dict())

    # ---- target (dqn.py:84-96) --------------------------------------------------------------------
    def _target_q(self, buffer: Any, indices: Any) -> torch.Tensor:
        batch = buffer[indices]  # batch.obs_next: s_{t+n}
        result = self(batch, input="obs_next")
        if self._target:  # target_Q = Q_old(s_, argmax(Q_new(s_, *)))
            target_q = self(batch, model="model_old", input="obs_next").logits
        else:
            target_q = result.logits
        if self._is_double:
            return target_q.gather(1, result.act.reshape(-1, 1)).squeeze(1)
        return target_q.max(dim=1)[0]

    def process_fn(self, batch: Batch, buffer: Any, indices: Any) -> Batch:
        return self.compute_nstep_return(batch, buffer, indices, self._target_q, self._gamma,
                                         self._n_step, self._rew_norm)

    def compute_q_value(self, logits: torch.Tensor, mask: Optional[Any]) -> torch.Tensor:
        if mask is not None:  # the masked q value should be smaller than logits.min()
            min_value = logits.min() - logits.max() - 1.0
            mask = torch.as_tensor(np.asarray(mask) if not isinstance(mask, torch.Tensor)
                                   else mask, device=logits.device)
            logits = logits + (1 - mask.to(logits.dtype)) * min_value
        return logits

    # ---- forward (dqn.py:120-170) -----------------------------------------------------------------
    def forward(self, batch: Batch, state: Any = None, model: str = "model",
                input: str = "obs", **kwargs: Any) -> Batch:
        net = getattr(self, model)
        obs = batch[input]
        obs_next = obs.obs if hasattr(obs, "obs") else obs
        logits, hidden = net(obs_next, state=state, info=batch.get("info"))
        q = self.compute_q_value(logits, getattr(obs, "mask", None))
        if not hasattr(self, "max_action_num"):
            self.max_action_num = q.shape[1]
        act = q.max(dim=1)[1]  # stays on the device (the reference pulls it to NumPy)
        return Batch(logits=logits, act=act, state=hidden)

    # ---- learn (dqn.py:172-194) -------------------------------------------------------------------
    def learn(self, batch: Batch, **kwargs: Any) -> Dict[str, float]:
        if self._target and self._iter % self._freq == 0:
            self.sync_weight()
        self.optim.zero_grad()
        weight = batch.pop("weight", 1.0)
        q = self(batch).logits
        act = batch.act
        if not isinstance(act, torch.Tensor):
            act = torch.as_tensor(np.asarray(act))
        q = q.gather(1, act.to(q.device).long().reshape(-1, 1)).squeeze(1)
        returns = batch.returns.flatten().to(q)
        td_error = returns - q
        if self._clip_loss_grad:
            loss = torch.nn.functional.huber_loss(q.reshape(-1, 1), returns.reshape(-1, 1),
                                                  reduction="mean")
        else:
            loss = (td_error.pow(2) * weight).mean()
        batch.weight = td_error.detach()  # prio-buffer
        loss.backward()
        self._sync_grads(self.model.parameters())
        self.optim.step()
        self._iter += 1
        return {"loss": loss.item()}

    def exploration_noise(self, act: Any, batch: Batch) -> Any:
        """Epsilon-greedy on the host RNG, as the reference draws it (dqn.py:196-210)."""
        if np.isclose(self.eps, 0.0):
            return act
        is_tensor = isinstance(act, torch.Tensor)
        a = act.cpu().numpy().copy() if is_tensor else act
        bsz = len(a)
        rand_mask = np.random.rand(bsz) < self.eps
        q = np.random.rand(bsz, self.max_action_num)  # [0, 1]
        if hasattr(batch.obs, "mask"):
            m = batch.obs.mask
            q += m.cpu().numpy() if isinstance(m, torch.Tensor) else m
        rand_act = q.argmax(axis=1)
        a[rand_mask] = rand_act[rand_mask]
        return torch.as_tensor(a, device=act.device) if is_tensor else a
