"""Synchronous advantage actor-critic on device-resident rollouts -- drop-in for
``tianshou.policy.A2CPolicy`` on the return path (tianshou/policy/modelfree/a2c.py:48-175).

``_compute_returns`` (a2c.py:83-117) never leaves the device: no ``.cpu().numpy()`` of the
critic outputs (a2c.py:91-93) and no ``to_torch_as`` copies back (a2c.py:115-116); GAE runs in
``csrc/gae.cu`` and the running return statistics are device scalars.
"""
from __future__ import annotations

from typing import Any, Callable, Dict, List, Optional

import numpy as np
import torch
from torch import nn

from ..data.batch import Batch
from . import _graph as G
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
        cuda_graph: bool = False,
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
        self._cuda_graph = cuda_graph  # this package's addition: graphed minibatch steps
        self._graphs: Optional["_GraphedLearner"] = None

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
    _graph_fields = ("adv", "returns")  # per-sample constants a minibatch step reads

    def _flat(self, t: torch.Tensor, n: int, what: str) -> torch.Tensor:
        t = t.detach().reshape(-1).float().contiguous()
        assert t.numel() == n, f"{what} must be one scalar per transition"
        return t

    def _minibatch_step(self, obs, act, mb, adv, returns) -> torch.Tensor:
        """a2c.py:125-152 for one minibatch; returns the four logged scalars (device)."""
        dist = self._dist(obs)[2]
        adv, ret = adv[mb], returns[mb]
        log_prob = dist.log_prob(act).reshape(len(adv), -1).transpose(0, 1)
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
        return torch.stack([loss.detach(), actor_loss.detach(), vf_loss.detach(),
                            ent_loss.detach()])

    def _learn_table(self, batch: Batch, batch_size: int, repeat: int) -> np.ndarray:
        """[minibatches, 4] logged scalars; eager loop or CUDA-graph replays."""
        if self._cuda_graph:
            if self._graphs is None:
                self._graphs = _GraphedLearner(self)
            return self._graphs.learn(batch, batch_size, repeat)
        return self._learn_eager(batch, batch_size, repeat)

    def _learn_eager(self, batch: Batch, batch_size: int, repeat: int) -> np.ndarray:
        n = len(batch)
        logged: List[torch.Tensor] = []
        for step in range(repeat):
            if getattr(self, "_recompute_adv", False) and step > 0:  # ppo.py:103-104
                batch = self._compute_returns(batch, self._buffer, self._indices)
            fields = [self._flat(batch[name], n, name) for name in self._graph_fields]
            for mb in _minibatches(n, batch_size, self.device):
                logged.append(self._minibatch_step(batch.obs[mb], batch.act[mb], mb, *fields))
        return torch.stack(logged).cpu().numpy() if logged else np.zeros((0, 4))  # one sync

    def learn(self, batch: Batch, batch_size: int, repeat: int, **kwargs: Any
              ) -> Dict[str, List[float]]:
        table = self._learn_table(batch, batch_size, repeat)
        return {
            "loss": table[:, 0].tolist(),
            "loss/actor": table[:, 1].tolist(),
            "loss/vf": table[:, 2].tolist(),
            "loss/ent": table[:, 3].tolist(),
        }


class _GraphedLearner:
    """``A2CPolicy.learn`` / ``PPOPolicy.learn`` as CUDA-graph replays.

    Static device state: the update's rows (``obs``, ``act``, ``logp_old``, ``adv``, ``returns``,
    ``v_s``), the concatenated permutations of all epochs, a cursor into them, and the loss
    log with its row counter.  A graph for minibatch size m reads ``perm[cursor : cursor + m]``,
    runs ``PPOPolicy._minibatch_step`` and advances cursor and row itself, so replays need no
    host input.  Graphs are keyed by (m, learning rates, train/eval mode) and dropped when a
    static buffer has to grow.  Warm-up iterations run for real on a side stream (allocator
    and lazy optimiser state), after which parameters, module buffers and optimiser state are
    restored in place, so capture does not change the training trajectory.
    """

    LOG_ROWS = 1 << 16

    def __init__(self, policy: "A2CPolicy") -> None:
        self.p = policy
        self.dev = policy.device
        self.static: Dict[str, torch.Tensor] = {}
        self.graphs: Dict[Any, torch.cuda.CUDAGraph] = {}
        G.prepare_optimizer(policy.optim, self.dev)
        self.cursor = torch.zeros(1, dtype=torch.int64, device=self.dev)
        self.row = torch.zeros(1, dtype=torch.int64, device=self.dev)
        # one spare row: warm-up iterations log there, so a capture in the middle of a learn
        # call (the merged tail's size) does not overwrite rows already written
        self.log = torch.zeros(self.LOG_ROWS + 1, 4, dtype=torch.float32, device=self.dev)

    # ---- static rows ---------------------------------------------------------------------------
    def _fit(self, key: str, src: torch.Tensor) -> None:
        cur = self.static.get(key)
        if (cur is None or cur.shape[0] < src.shape[0] or cur.shape[1:] != src.shape[1:]
                or cur.dtype != src.dtype):
            self.static[key] = torch.empty_like(src, device=self.dev)
            self.graphs.clear()  # captured addresses are stale
            cur = self.static[key]
        cur[:src.shape[0]].copy_(src)

    def _load_fields(self, batch: Batch, n: int) -> None:
        p = self.p
        for name in p._graph_fields:
            self._fit(name, p._flat(batch[name], n, name))

    # ---- capture -------------------------------------------------------------------------------
    def _body(self, m: int) -> None:
        s, p = self.static, self.p
        mb = s["perm"].index_select(0, self.cursor + s["arange"][:m])
        with G.no_distribution_validation():
            vals = p._minibatch_step(
                s["obs"].index_select(0, mb), s["act"].index_select(0, mb), mb,
                *[s[name] for name in p._graph_fields])
        self.log.index_copy_(0, self.row, vals[None])
        self.cursor += m
        self.row += 1

    def _guarded(self) -> List[torch.Tensor]:
        p = self.p
        return (list(p._params) + list(p.actor.buffers()) + list(p.critic.buffers())
                + [self.cursor, self.row])

    def _warm_row(self) -> None:
        self.cursor.zero_()
        self.row.fill_(self.LOG_ROWS)  # warm-up iterations log into the spare row

    def _graph(self, m: int) -> torch.cuda.CUDAGraph:
        key = (m, G.lr_key(self.p.optim), self.p.training)
        g = self.graphs.get(key)
        if g is None:
            g = self.graphs[key] = G.capture(lambda: self._body(m), self._guarded,
                                             self.p.optim, self.dev, self._warm_row)
        return g

    # ---- learn ---------------------------------------------------------------------------------
    def learn(self, batch: Batch, batch_size: int, repeat: int) -> np.ndarray:
        p = self.p
        n = len(batch)
        if not isinstance(batch.obs, torch.Tensor):
            raise NotImplementedError("cuda_graph=True needs tensor observations")
        bounds = _split_bounds(n, batch_size, merge_last=True)
        if repeat <= 0 or not bounds:
            return np.zeros((0, 4))
        self._fit("obs", batch.obs)
        self._fit("act", batch.act)
        self._load_fields(batch, n)
        perms = np.concatenate([np.random.permutation(n) for _ in range(repeat)])  # batch.py:905
        self._fit("perm", torch.from_numpy(perms).to(self.dev))
        m_max = max(hi - lo for lo, hi in bounds)
        if "arange" not in self.static or self.static["arange"].numel() < m_max:
            self.static["arange"] = torch.arange(m_max, dtype=torch.int64, device=self.dev)
            self.graphs.clear()
        self.cursor.zero_()
        self.row.zero_()
        tables: List[np.ndarray] = []
        rows = 0
        for step in range(repeat):
            if getattr(p, "_recompute_adv", False) and step > 0:
                batch = p._compute_returns(batch, p._buffer, p._indices)
                self._load_fields(batch, n)
            for lo, hi in bounds:
                if rows == self.LOG_ROWS:  # log full: read it out and start over
                    tables.append(self.log[:rows].cpu().numpy().copy())
                    self.row.zero_()
                    rows = 0
                self._graph(hi - lo).replay()
                rows += 1
        tables.append(self.log[:rows].cpu().numpy())  # the one sync of a normal-sized learn
        return np.concatenate(tables) if len(tables) > 1 else tables[0]
