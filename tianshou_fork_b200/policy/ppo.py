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
* ``cuda_graph=True`` (opt-in): one minibatch step -- index gather, actor and
  critic forward, fused loss, backward, gradient clipping, optimiser step, loss logging -- is
  captured once per minibatch size into a CUDA graph that advances its own cursor into the
  epoch's permutation on the device, so ``learn`` is one graph launch per minibatch and one
  host sync per call.  The arithmetic is the eager path's, kernel for kernel.
"""
from __future__ import annotations

from typing import Any, Callable, Dict, List, Optional, Tuple

import numpy as np
import torch
from torch import nn

from .. import _lib as L
from ..data.batch import Batch
from .a2c import A2CPolicy, _GraphedLearner
from .pg import RunningMeanStd, _minibatches, _split_bounds  # noqa: F401 (re-exported)

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
    sh = L.shim()
    if sh is not None:  # the same C-ABI call from C++ (csrc/shim.cpp): no ctypes, C++ autograd node
        return tuple(sh.ppo_loss(logp, value, logp_old, adv, returns, v_s, mb_index,
                                 _ws(logp.device), float(eps_clip), float(dual_clip or 0.0),
                                 bool(value_clip), bool(norm_adv), float(eps)))
    return _PPOLoss.apply(logp, value, logp_old, adv, returns, v_s, mb_index, eps_clip,
                          dual_clip, value_clip, norm_adv, eps)


class PPOPolicy(A2CPolicy):
    """PPO (arXiv:1707.06347) with GAE on the device and the fused clip loss.

    Constructor arguments follow the reference (ppo.py:63-86, a2c.py:48-75, pg.py:49-85):
    ``actor(obs, state=, info=) -> (logits, state)``, ``critic(obs) -> value``,
    ``dist_fn(*logits)``.  ``cuda_graph`` is this package's addition (see the module docstring).
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
        cuda_graph: bool = False,
        **kwargs: Any,
    ) -> None:
        super().__init__(actor, critic, optim, dist_fn, cuda_graph=cuda_graph, **kwargs)
        self._eps_clip = eps_clip
        assert dual_clip is None or dual_clip > 1.0, \
            "Dual-clip PPO parameter should greater than 1.0."
        self._dual_clip = dual_clip
        self._value_clip = value_clip
        self._norm_adv = advantage_normalization
        self._recompute_adv = recompute_advantage

    _graph_fields = ("logp_old", "adv", "returns", "v_s")  # per-sample constants of a step

    def process_fn(self, batch: Batch, buffer: Any, indices: Any) -> Batch:
        if self._recompute_adv:  # ppo.py:90-92
            self._buffer, self._indices = buffer, indices
        batch = self._compute_returns(batch, buffer, indices)
        batch.act = self._act_tensor(batch)
        with torch.no_grad():
            batch.logp_old = self._dist(batch.obs)[2].log_prob(batch.act)  # ppo.py:95-96
        return batch

    def _minibatch_step(self, obs, act, mb, logp_old, adv, returns, v_s) -> torch.Tensor:
        """ppo.py:106-151 for one minibatch; returns the four logged scalars (device)."""
        dist = self._dist(obs)[2]
        logp = dist.log_prob(act)
        value = self.critic(obs).flatten()
        clip_loss, vf_loss, _ = ppo_loss(
            logp, value, logp_old, adv, returns, v_s, mb, self._eps_clip,
            self._dual_clip, self._value_clip, self._norm_adv, self._eps)
        ent_loss = dist.entropy().mean()
        loss = clip_loss + self._weight_vf * vf_loss - self._weight_ent * ent_loss
        self.optim.zero_grad()
        loss.backward()
        self._sync_grads(self._params)
        if self._grad_norm:
            nn.utils.clip_grad_norm_(self._params, max_norm=self._grad_norm)
        self.optim.step()
        return torch.stack([loss.detach(), clip_loss.detach(), vf_loss.detach(),
                            ent_loss.detach()])

    # ---- learn (ppo.py:99-162) ------------------------------------------------------------------
    def learn(self, batch: Batch, batch_size: int, repeat: int, **kwargs: Any
              ) -> Dict[str, List[float]]:
        table = self._learn_table(batch, batch_size, repeat)
        return {
            "loss": table[:, 0].tolist(),
            "loss/clip": table[:, 1].tolist(),
            "loss/vf": table[:, 2].tolist(),
            "loss/ent": table[:, 3].tolist(),
        }
