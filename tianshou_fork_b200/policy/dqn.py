"""Deep Q-learning on a device-resident (prioritized) replay buffer -- drop-in for
``tianshou.policy.DQNPolicy`` on the n-step / PER path (tianshou/policy/modelfree/dqn.py:40-215).

The pipeline of one ``update`` never leaves the device: ``buffer.sample`` (PER descent +
importance weights, ``csrc/segtree.cu``) -> ``process_fn`` = ``compute_nstep_return``
(``csrc/nstep.cu``, the target network evaluated on ``buffer[terminal]``) -> ``learn`` ->
``post_process_fn`` = ``buffer.update_weight(indices, td_error)`` (policy/base.py:285-286).
The only host read is the logged ``loss.item()`` the reference API returns (dqn.py:194).

``cuda_graph=True`` (opt-in) captures that whole pipeline for a given
(buffer, sample size) once and replays it per ``update``: the host draws the uniforms (PER) or
the indices (uniform replay) from ``np.random`` exactly as the reference does, uploads them
into a static tensor, refreshes the buffer's ring bookkeeping in place and launches one graph.
"""
from __future__ import annotations

from copy import deepcopy
from typing import Any, Dict, Optional

import numpy as np
import torch

from ..data.batch import Batch
from . import _graph as G
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
        cuda_graph: bool = False,
        **kwargs: Any,
    ) -> None:
        super().__init__(**kwargs)
        self._cuda_graph = cuda_graph
        self._graphed: Dict[Any, "_GraphedUpdate"] = {}
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
        if hasattr(buffer, "obs_next_of"):  # only obs_next (s_{t+n}) and info are read below
            batch = Batch(obs_next=buffer.obs_next_of(indices),
                          info=buffer.get(indices, "info", Batch()))
        else:
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
    def _learn_step(self, batch: Batch) -> torch.Tensor:
        """dqn.py:175-192 without the target sync, the iteration count and the host read."""
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
        return loss.detach()

    def learn(self, batch: Batch, **kwargs: Any) -> Dict[str, float]:
        if self._target and self._iter % self._freq == 0:
            self.sync_weight()
        loss = self._learn_step(batch)
        self._iter += 1
        return {"loss": loss.item()}

    def update(self, sample_size: int, buffer: Optional[Any], **kwargs: Any) -> Dict[str, Any]:
        if not self._cuda_graph or buffer is None or sample_size <= 0 or len(buffer) == 0:
            return super().update(sample_size, buffer, **kwargs)
        key = (id(buffer), int(sample_size))
        g = self._graphed.get(key)
        if g is None:
            g = self._graphed[key] = _GraphedUpdate(self, buffer, int(sample_size))
        return g.update()


class _GraphedUpdate:
    """``DQNPolicy.update(sample_size, buffer)`` as one CUDA-graph replay.

    Static inputs: the step's uniforms (prioritized buffers: ``np.random.rand``, prio.py:61) or
    its indices (uniform buffers: the buffer's own host draw).  Body: PER descent + importance
    weights -> ``buffer[idx]`` gathers -> n-step return with the target network on
    ``buffer[terminal]`` -> loss, backward, optimiser step -> ``update_weight``.  Everything
    the body reads of the buffer's bookkeeping sits at fixed device addresses that ``_state()``
    refreshes in place, so transitions added between updates are seen by the next replay.
    Graphs are keyed by learning rates, beta, train/eval mode and whether the update is one
    that copies the online weights into the target network (the reference does that inside
    ``learn``, i.e. after the targets were computed with the old target network).
    """

    def __init__(self, policy: DQNPolicy, buffer: Any, bsz: int) -> None:
        self.p, self.buf, self.bsz = policy, buffer, bsz
        self.dev = policy.device
        self.per = hasattr(buffer, "sample_with_uniforms")
        G.prepare_optimizer(policy.optim, self.dev)
        self.u = torch.zeros(bsz, dtype=torch.float64, device=self.dev)
        self.idx = torch.zeros(bsz, dtype=torch.int64, device=self.dev)
        self.loss = torch.zeros(1, dtype=torch.float32, device=self.dev)
        self.last_indices = torch.zeros(bsz, dtype=torch.int64, device=self.dev)
        self.graphs: Dict[Any, torch.cuda.CUDAGraph] = {}

    def _draw(self) -> None:
        if self.per and not self.buf.device_rng:
            self.u.copy_(torch.from_numpy(np.random.rand(self.bsz)))
        elif self.per:
            self.u.copy_(torch.rand(self.bsz, dtype=torch.float64, device=self.dev))
        else:
            self.idx.copy_(self.buf.sample_indices(self.bsz))

    def _body(self, sync: bool) -> None:
        p, buf = self.p, self.buf
        idx = buf.sample_with_uniforms(self.u) if self.per else self.idx
        self.last_indices.copy_(idx)  # what this update sampled, for callers that log it
        batch = buf[idx]
        batch = p.process_fn(batch, buf, idx)
        if sync:  # the reference syncs inside learn(): AFTER the targets, BEFORE the step
            p.sync_weight()
        self.loss.copy_(p._learn_step(batch).reshape(1))
        p.post_process_fn(batch, buf, idx)

    def _guarded(self):
        p, buf = self.p, self.buf
        out = list(p.model.parameters()) + list(p.model.buffers())
        if p._target:  # the warm-up of the syncing variant overwrites the target network
            out += list(p.model_old.parameters()) + list(p.model_old.buffers())
        if self.per:  # the warm-up iterations really update priorities
            out += [buf.weight._value, buf._prio]
        return out

    def _key(self) -> tuple:
        buf = self.buf  # storage identity: a re-allocated field invalidates captured addresses
        return (G.lr_key(self.p.optim), getattr(buf, "_beta", None), self.p.training,
                tuple(v.data_ptr() if isinstance(v, torch.Tensor) else id(v)
                      for v in buf._meta.values()))

    def update(self) -> Dict[str, float]:
        p, buf = self.p, self.buf
        p.updating = True
        sync = bool(p._target and p._iter % p._freq == 0)  # dqn.py:173-174
        buf._state()  # ring bookkeeping of transitions added since the last update, in place
        key = self._key() + (sync,)
        graph = self.graphs.get(key)
        if graph is None:
            self._draw_for_warmup()
            graph = self.graphs[key] = G.capture(lambda: self._body(sync), self._guarded,
                                                 p.optim, self.dev)
        self._draw()
        graph.replay()
        if self.per:  # the replayed update_weight left the top of the sum tree to the next reader
            buf.weight._top_stale = True
        p._iter += 1
        if p.lr_scheduler is not None:
            p.lr_scheduler.step()
        p.updating = False
        return {"loss": self.loss.item()}

    def _draw_for_warmup(self) -> None:
        """Warm-up inputs must not consume the host RNG stream the real updates draw from."""
        state = np.random.get_state()
        self._draw()
        np.random.set_state(state)
