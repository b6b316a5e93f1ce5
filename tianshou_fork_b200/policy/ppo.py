"""PPO on device-resident rollouts with the fused loss kernel -- drop-in for
``tianshou.policy.PPOPolicy`` on the hot path (tianshou/policy/modelfree/ppo.py:63-162 with
the A2C / PG glue it inherits: a2c.py:48-117, pg.py:49-120).

What changes relative to the reference:
* ``_compute_returns`` never leaves the device: no ``.cpu().numpy()`` (a2c.py:91-93) and no
  ``to_torch_as`` copies back (a2c.py:115-116); GAE runs in ``csrc/gae.cu``.
* ``learn`` replaces the ~25 eager elementwise/reduction launches per minibatch
  (ppo.py:109-139) by one launch of ``csrc/ppo_loss.cu`` (forward values + gradients), gathers
  the per-sample constants through the minibatch index inside the kernel, and collects the
  four logged scalars on the device: one host sync per ``learn`` call instead of four per
  minibatch (ppo.py:152-155).
* minibatch indices come from the host ``np.random.permutation`` exactly as ``Batch.split``
  draws them (data/batch.py:896-912), so a seeded run visits the same minibatches.
"""
from __future__ import annotations

from typing import Any, Callable, Dict, List, Optional, Tuple

import numpy as np
import torch
from torch import nn

from .. import _lib as L
from ..data.batch import Batch
from .base import BasePolicy

_loss_ws: Dict[int, torch.Tensor] = {}


def _ws(device: torch.device) -> torch.Tensor:
    key = device.index or 0
    if key not in _loss_ws:
        n = L.lib().ts_ppo_loss_workspace_bytes(1)
        _loss_ws[key] = torch.zeros(n, dtype=torch.uint8, device=device)
    return _loss_ws[key]


class _PPOLoss(torch.autograd.Function):
    """(clip_loss, vf_loss) of ppo.py:109-139 as one kernel; backward is two scalings."""

    @staticmethod
    def forward(ctx, logp, value, logp_old, adv, returns, v_s, mb_index, eps_clip, dual_clip,
                value_clip, norm_adv, eps):
        m = logp.numel()
        dev = logp.device
        f32 = lambda t: t.detach().reshape(-1).to(torch.float32).contiguous()
        lp, val = f32(logp), f32(value)
        assert val.numel() == m
        out = torch.empty(4, dtype=torch.float32, device=dev)
        dlogp = torch.empty(m, dtype=torch.float32, device=dev)
        dvalue = torch.empty(m, dtype=torch.float32, device=dev)
        ws = _ws(dev)
        L.check(L.lib().ts_ppo_loss(
            L.ptr(lp), L.ptr(val), L.ptr(logp_old), L.ptr(adv), L.ptr(returns),
            L.ptr(v_s) if v_s is not None else None,
            L.ptr(mb_index) if mb_index is not None else None, m, float(eps_clip),
            float(dual_clip or 0.0), int(bool(value_clip)), int(bool(norm_adv)), float(eps),
            L.ptr(out), L.ptr(dlogp), L.ptr(dvalue), L.ptr(ws), ws.numel(), L.stream()),
            "ts_ppo_loss")
        ctx.save_for_backward(dlogp, dvalue)
        ctx.shapes = (logp.shape, value.shape, logp.dtype, value.dtype)
        return out[0], out[1], out[2:4]

    @staticmethod
    def backward(ctx, g_clip, g_vf, _g_stats):
        dlogp, dvalue = ctx.saved_tensors
        ls, vs, ld, vd = ctx.shapes
        g_logp = (dlogp * g_clip).reshape(ls).to(ld)
        g_value = (dvalue * g_vf).reshape(vs).to(vd)
        return (g_logp, g_value) + (None,) * 10


def ppo_loss(logp: torch.Tensor, value: torch.Tensor, logp_old: torch.Tensor,
             adv: torch.Tensor, returns: torch.Tensor, v_s: Optional[torch.Tensor] = None,
             mb_index: Optional[torch.Tensor] = None, eps_clip: float = 0.2,
             dual_clip: Optional[float] = None, value_clip: bool = False,
             norm_adv: bool = True, eps: float = 1e-8
             ) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """Fused PPO losses.  ``logp``/``value`` are the minibatch's network outputs (length m);
    ``logp_old``/``adv``/``returns``/``v_s`` are f32 CUDA arrays indexed by ``mb_index`` (or of
    length m when ``mb_index`` is None).  Returns ``(clip_loss, vf_loss, [adv_mean, adv_std])``;
    gradients flow to ``logp`` and ``value`` only, like the reference where adv/returns/
    logp_old are constants."""
    for t in (logp_old, adv, returns) + ((v_s,) if v_s is not None else ()):
        assert t.is_cuda and t.dtype == torch.float32 and t.is_contiguous()
    if mb_index is not None:
        assert mb_index.dtype == torch.int64 and mb_index.is_contiguous()
        assert mb_index.numel() == logp.numel()
    assert not value_clip or v_s is not None
    return _PPOLoss.apply(logp, value, logp_old, adv, returns, v_s, mb_index, eps_clip,
                          dual_clip, value_clip, norm_adv, eps)


class RunningMeanStd:
    """Running mean / variance of a stream (tianshou/utils/statistics.py:69-114) kept in
    device scalars so that return normalisation (a2c.py:98-112) does not sync."""

    def __init__(self, device: torch.device, eps: float = np.finfo(np.float32).eps.item()):
        self.mean = torch.zeros((), dtype=torch.float64, device=device)
        self.var = torch.ones((), dtype=torch.float64, device=device)
        self.count = 0
        self.eps = eps

    def update(self, x: torch.Tensor) -> None:
        x = x.to(torch.float64)
        batch_mean, batch_var = x.mean(), x.var(unbiased=False)
        batch_count = x.numel()
        delta = batch_mean - self.mean
        total = self.count + batch_count
        new_mean = self.mean + delta * batch_count / total
        m_a, m_b = self.var * self.count, batch_var * batch_count
        m_2 = m_a + m_b + delta ** 2 * self.count * batch_count / total
        self.mean, self.var, self.count = new_mean, m_2 / total, total


class PPOPolicy(BasePolicy):
    """PPO (arXiv:1707.06347) with GAE on the device and the fused clip loss.

    Constructor arguments follow the reference (ppo.py:63-86, a2c.py:48-75, pg.py:49-85):
    ``actor(obs) -> (logits, state)``, ``critic(obs) -> value``, ``dist_fn(*logits)``.
    """

    def __init__(
        self,
        actor: nn.Module,
        critic: nn.Module,
        optim: torch.optim.Optimizer,
        dist_fn: Callable[..., torch.distributions.Distribution],
        eps_clip: float = 0.2,
        dual_clip: Optional[float] = None,
        value_clip: bool = False,
        advantage_normalization: bool = True,
        recompute_advantage: bool = False,
        vf_coef: float = 0.5,
        ent_coef: float = 0.01,
        max_grad_norm: Optional[float] = None,
        gae_lambda: float = 0.95,
        max_batchsize: int = 256,
        discount_factor: float = 0.99,
        reward_normalization: bool = False,
        lr_scheduler: Optional[Any] = None,
    ) -> None:
        super().__init__()
        self.actor, self.critic, self.optim, self.dist_fn = actor, critic, optim, dist_fn
        assert dual_clip is None or dual_clip > 1.0, \
            "Dual-clip PPO parameter should greater than 1.0."
        assert 0.0 <= gae_lambda <= 1.0, "GAE lambda should be in [0, 1]."
        assert 0.0 <= discount_factor <= 1.0, "discount factor should be in [0, 1]"
        self._eps_clip, self._dual_clip, self._value_clip = eps_clip, dual_clip, value_clip
        self._norm_adv, self._recompute_adv = advantage_normalization, recompute_advantage
        self._weight_vf, self._weight_ent, self._grad_norm = vf_coef, ent_coef, max_grad_norm
        self._lambda, self._batch, self._gamma = gae_lambda, max_batchsize, discount_factor
        self._rew_norm = reward_normalization
        self._eps = 1e-8
        self.lr_scheduler = lr_scheduler
        self.device = next(actor.parameters()).device
        self.ret_rms = RunningMeanStd(self.device)
        self._params = list(actor.parameters()) + [
            p for p in critic.parameters() if all(p is not q for q in actor.parameters())]

    # ---- forward (pg.py:122-160) ------------------------------------------------------------------
    def __call__(self, batch: Batch, state: Any = None) -> Batch:
        logits, hidden = self.actor(batch.obs, state=state)
        dist = self.dist_fn(*logits) if isinstance(logits, tuple) else self.dist_fn(logits)
        return Batch(logits=logits, state=hidden, dist=dist)

    # ---- returns (a2c.py:83-117) ------------------------------------------------------------------
    def _values(self, obs: torch.Tensor) -> torch.Tensor:
        n = obs.shape[0]
        # Batch.split(self._batch, shuffle=False, merge_last=True)  (a2c.py:88)
        out = [self.critic(obs[lo:hi]) for lo, hi in _split_bounds(n, self._batch, True)]
        return torch.cat(out, dim=0).flatten()

    def _compute_returns(self, batch: Batch, buffer: Any, indices: torch.Tensor) -> Batch:
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

    def process_fn(self, batch: Batch, buffer: Any, indices: torch.Tensor) -> Batch:
        if self._recompute_adv:
            self._buffer, self._indices = buffer, indices
        batch = self._compute_returns(batch, buffer, indices)
        act = batch.act
        if not isinstance(act, torch.Tensor):
            act = torch.as_tensor(np.asarray(act), device=self.device)
        batch.act = act.to(self.device)
        with torch.no_grad():
            batch.logp_old = self._log_prob(self(batch).dist, batch.act)
        return batch

    @staticmethod
    def _log_prob(dist, act: torch.Tensor) -> torch.Tensor:
        return dist.log_prob(act)

    # ---- learn (ppo.py:99-162) ------------------------------------------------------------------
    def learn(self, batch: Batch, batch_size: int, repeat: int, **kwargs: Any
              ) -> Dict[str, List[float]]:
        n = len(batch)
        logged: List[torch.Tensor] = []
        for step in range(repeat):
            if self._recompute_adv and step > 0:
                batch = self._compute_returns(batch, self._buffer, self._indices)
            logp_old = batch.logp_old.detach().reshape(-1).float().contiguous()
            adv = batch.adv.detach().reshape(-1).float().contiguous()
            returns = batch.returns.detach().reshape(-1).float().contiguous()
            v_s = batch.v_s.detach().reshape(-1).float().contiguous()
            assert logp_old.numel() == n, "log_prob must be one scalar per transition"
            perm = torch.from_numpy(np.random.permutation(n)).to(self.device)  # batch.py:905
            for lo, hi in _split_bounds(n, batch_size, merge_last=True):
                mb = perm[lo:hi]
                obs, act = batch.obs[mb], batch.act[mb]
                dist = self(Batch(obs=obs)).dist
                logp = self._log_prob(dist, act)
                value = self.critic(obs).flatten()
                clip_loss, vf_loss, _ = ppo_loss(
                    logp, value, logp_old, adv, returns, v_s, mb, self._eps_clip,
                    self._dual_clip, self._value_clip, self._norm_adv, self._eps)
                ent_loss = dist.entropy().mean()
                loss = clip_loss + self._weight_vf * vf_loss - self._weight_ent * ent_loss
                self.optim.zero_grad()
                loss.backward()
                if self._grad_norm:
                    nn.utils.clip_grad_norm_(self._params, max_norm=self._grad_norm)
                self.optim.step()
                logged.append(torch.stack([loss.detach(), clip_loss.detach(),
                                           vf_loss.detach(), ent_loss.detach()]))
        table = torch.stack(logged).cpu().numpy() if logged else np.zeros((0, 4))  # one sync
        return {
            "loss": table[:, 0].tolist(),
            "loss/clip": table[:, 1].tolist(),
            "loss/vf": table[:, 2].tolist(),
            "loss/ent": table[:, 3].tolist(),
        }


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
