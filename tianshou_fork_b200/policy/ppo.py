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
from . import _graph as G
from ..data.batch import Batch
from .a2c import A2CPolicy
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
        super().__init__(actor, critic, optim, dist_fn, **kwargs)
        self._eps_clip = eps_clip
        assert dual_clip is None or dual_clip > 1.0, \
            "Dual-clip PPO parameter should greater than 1.0."
        self._dual_clip = dual_clip
        self._value_clip = value_clip
        self._norm_adv = advantage_normalization
        self._recompute_adv = recompute_advantage
        self._cuda_graph = cuda_graph
        self._graphs: Optional[_GraphedLearner] = None

    def process_fn(self, batch: Batch, buffer: Any, indices: Any) -> Batch:
        if self._recompute_adv:  # ppo.py:90-92
            self._buffer, self._indices = buffer, indices
        batch = self._compute_returns(batch, buffer, indices)
        batch.act = self._act_tensor(batch)
        with torch.no_grad():
            batch.logp_old = self._dist(batch.obs)[2].log_prob(batch.act)  # ppo.py:95-96
        return batch

    def _flat(self, t: torch.Tensor, n: int, what: str) -> torch.Tensor:
        t = t.detach().reshape(-1).float().contiguous()
        assert t.numel() == n, f"{what} must be one scalar per transition"
        return t

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
        if self._cuda_graph:
            if self._graphs is None:
                self._graphs = _GraphedLearner(self)
            table = self._graphs.learn(batch, batch_size, repeat)
        else:
            table = self._learn_eager(batch, batch_size, repeat)
        return {
            "loss": table[:, 0].tolist(),
            "loss/clip": table[:, 1].tolist(),
            "loss/vf": table[:, 2].tolist(),
            "loss/ent": table[:, 3].tolist(),
        }

    def _learn_eager(self, batch: Batch, batch_size: int, repeat: int) -> np.ndarray:
        n = len(batch)
        logged: List[torch.Tensor] = []
        for step in range(repeat):
            if self._recompute_adv and step > 0:
                batch = self._compute_returns(batch, self._buffer, self._indices)
            logp_old = self._flat(batch.logp_old, n, "log_prob")
            adv = self._flat(batch.adv, n, "adv")
            returns = self._flat(batch.returns, n, "returns")
            v_s = self._flat(batch.v_s, n, "v_s")
            for mb in _minibatches(n, batch_size, self.device):
                logged.append(self._minibatch_step(batch.obs[mb], batch.act[mb], mb, logp_old,
                                                   adv, returns, v_s))
        return torch.stack(logged).cpu().numpy() if logged else np.zeros((0, 4))  # one sync


class _GraphedLearner:
    """``PPOPolicy.learn`` as CUDA-graph replays.

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

    def __init__(self, policy: PPOPolicy) -> None:
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

    def _load_returns(self, batch: Batch, n: int) -> None:
        p = self.p
        self._fit("adv", p._flat(batch.adv, n, "adv"))
        self._fit("returns", p._flat(batch.returns, n, "returns"))
        self._fit("v_s", p._flat(batch.v_s, n, "v_s"))

    # ---- capture -------------------------------------------------------------------------------
    def _body(self, m: int) -> None:
        s, p = self.static, self.p
        mb = s["perm"].index_select(0, self.cursor + s["arange"][:m])
        with G.no_distribution_validation():
            vals = p._minibatch_step(
                s["obs"].index_select(0, mb), s["act"].index_select(0, mb), mb,
                s["logp_old"], s["adv"], s["returns"], s["v_s"])
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
        self._fit("logp_old", p._flat(batch.logp_old, n, "log_prob"))
        self._load_returns(batch, n)
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
            if p._recompute_adv and step > 0:
                batch = p._compute_returns(batch, p._buffer, p._indices)
                self._load_returns(batch, n)
            for lo, hi in bounds:
                if rows == self.LOG_ROWS:  # log full: read it out and start over
                    tables.append(self.log[:rows].cpu().numpy().copy())
                    self.row.zero_()
                    rows = 0
                self._graph(hi - lo).replay()
                rows += 1
        tables.append(self.log[:rows].cpu().numpy())  # the one sync of a normal-sized learn
        return np.concatenate(tables) if len(tables) > 1 else tables[0]
