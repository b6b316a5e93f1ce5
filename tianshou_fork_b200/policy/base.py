"""Return estimators on the device -- drop-ins for the static methods of
``tianshou.policy.BasePolicy`` (tianshou/policy/base.py:317-524).

``compute_episodic_return`` and ``compute_nstep_return`` keep the reference signatures and
semantics; the numba kernels ``_gae_return`` / ``_nstep_return`` and the NumPy glue around
them are replaced by the CUDA kernels in ``csrc/gae.cu`` / ``csrc/nstep.cu``.

Array kinds: the reference takes and returns NumPy.  Here NumPy inputs give NumPy outputs
(f64, as numba produces) and CUDA-tensor inputs give CUDA-tensor outputs with no host
synchronisation; ``buffer`` may be one of this package's device-resident buffers or any
object exposing the reference buffer attributes as NumPy arrays (they are then staged to the
device -- there is no CPU compute path).
"""
from __future__ import annotations

from typing import Any, Callable, Dict, Optional, Tuple, Union

import numpy as np
import torch
from torch import nn

from .. import _lib as L
from ..data.batch import Batch
from ..data.buffer import ReplayBuffer

ArrayLike = Union[np.ndarray, torch.Tensor]

_ws_cache: Dict[Tuple[int, int], torch.Tensor] = {}

# From this many transitions on, a batch that is the buffer's own listing goes through
# ts_mark_unfinished + the flat kernel without indices (saves 9 of 30 bytes per transition;
# below it the two extra launches cost more host time than the kernel saves: 131 072
# transitions 48 vs 29 us, 8.4 M transitions 47.6 vs 66.0 us).
LISTING_MIN_N = 1 << 21

# NumPy callers of compute_episodic_return on a canonical rollout of at least this many
# transitions go through the chunked host pipeline (pipeline.GaeHostCall): copies of successive
# row chunks overlap with the kernel and with each other instead of running back to back.
HOST_CALL_MIN_N = 1 << 18


def _device_of(buffer: Any) -> torch.device:
    dev = getattr(buffer, "device", None)
    if isinstance(dev, torch.device) and dev.type == "cuda":
        return dev
    return torch.device(f"cuda:{torch.cuda.current_device()}")


def _dev(x: Any, device: torch.device, dtype: Optional[torch.dtype] = None) -> torch.Tensor:
    if isinstance(x, torch.Tensor):
        if (x.device == device and (dtype is None or x.dtype == dtype) and not x.requires_grad
                and x.is_contiguous()):
            return x  # nothing to do: every no-op torch call below still costs ~1-2 us of host time
        t = x.detach().to(device)
    else:
        t = torch.from_numpy(np.ascontiguousarray(np.asarray(x))).to(device)
    if dtype is not None and t.dtype != dtype:
        t = t.to(dtype)
    return t.contiguous()


def _flat(t: torch.Tensor) -> torch.Tensor:
    return t if t.ndim == 1 else t.reshape(-1)


def _float_dev(x: Any, device: torch.device) -> torch.Tensor:
    t = _dev(x, device)
    if t.dtype not in (torch.float32, torch.float64):
        t = t.to(torch.float64)
    return t if t.ndim == 1 else t.reshape(-1).contiguous()


def _workspace(nbytes: int, device: torch.device) -> torch.Tensor:
    key = (device.index or 0, L.stream())
    ws = _ws_cache.get(key)
    if ws is None or ws.numel() < nbytes:
        # zero-initialised once: the flat kernel keeps its call count in it (include/ts_b200.h)
        ws = torch.zeros(max(nbytes, 1 << 16), dtype=torch.uint8, device=device)
        _ws_cache[key] = ws
    return ws


def _buffer_layout(buffer: Any, device: torch.device):
    """Device copies of (offset[B+1], last_index[B], lengths[B], B, sub_size, done, terminated)."""
    if isinstance(buffer, ReplayBuffer):
        st = buffer._state()
        return (st["offset"], st["last_index"], st["lengths"], st["B"], st["sub_size"],
                L.as_u8(buffer._meta["done"]), L.as_u8(buffer._meta["terminated"]))
    # a reference-style host buffer: stage the fields the kernels read
    if hasattr(buffer, "_extend_offset"):
        off = np.asarray(buffer._extend_offset, dtype=np.int64)
        last = np.asarray(buffer.last_index, dtype=np.int64)
        lens = np.asarray(buffer._lengths, dtype=np.int64)
        sizes = np.diff(off)
        sub = int(sizes[0]) if np.all(sizes == sizes[0]) else 0
    else:
        size = max(len(buffer), 1)
        off = np.array([0, size], dtype=np.int64)
        last = np.asarray(buffer.last_index, dtype=np.int64).reshape(1)
        lens = np.array([len(buffer)], dtype=np.int64)
        sub = size
    return (_dev(off, device), _dev(last, device), _dev(lens, device), last.size, sub,
            L.as_u8(_dev(buffer.done, device, torch.bool)),
            L.as_u8(_dev(buffer.terminated, device, torch.bool)))


def _episodic_return_host_call(batch: Batch, dev: torch.device, canon: Tuple[int, int], v_s_: Any,
                               v_s: Any, gamma: float, gae_lambda: float,
                               out_dtype: Optional[torch.dtype]):
    """The reference's calling convention at rollout scale: critic values (and possibly the
    batch fields) as host arrays in, NumPy out (policy/base.py:337-346).  Canonical [E, T]
    batches only; returns None when a dtype falls outside what the rows kernel takes."""
    from ..pipeline import gae_host_call
    E, T = canon
    n = E * T

    def float_dtype(x):
        d = x.dtype
        if isinstance(x, torch.Tensor):
            return d if d in (torch.float32, torch.float64) else None
        return {np.dtype(np.float32): torch.float32,
                np.dtype(np.float64): torch.float64}.get(np.dtype(d))

    rew = batch.rew
    vs, vn = np.asarray(v_s) if not isinstance(v_s, torch.Tensor) else v_s, \
        np.asarray(v_s_) if not isinstance(v_s_, torch.Tensor) else v_s_
    rd, vd, nd = float_dtype(rew), float_dtype(vs), float_dtype(vn)
    if rd is None or vd is None or nd is None or any(
            int(np.prod(x.shape)) != n for x in (rew, vs, vn, batch.terminated, batch.truncated)):
        return None
    val_dtype = vd if vd == nd else torch.float64  # one value dtype per launch: widen
    ret, adv = gae_host_call(dev).run(
        E, T, rew, vs, vn, batch.terminated, batch.truncated, gamma, gae_lambda, rd, val_dtype,
        out_dtype if out_dtype is not None else torch.float64)
    return ret, adv


def allreduce_mean_grads(params: Any, group: Any = None) -> None:
    """Average ``p.grad`` over the process group with ONE collective: the gradients of the
    path's networks are a few hundred KB, so they travel as a single flat bucket (NCCL over
    NVLink on GPUs, capturable in a CUDA graph; gloo in the CPU tests)."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        raise RuntimeError("set_data_parallel() needs an initialised torch.distributed group")
    world = dist.get_world_size(group)
    grads = [p.grad for p in params if p.grad is not None]
    if world == 1 or not grads:
        return
    flat = torch.cat([g.reshape(-1) for g in grads])
    dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    flat.div_(world)
    offset = 0
    for g in grads:
        n = g.numel()
        g.copy_(flat[offset:offset + n].view_as(g))
        offset += n


class BasePolicy(nn.Module):
    """Static return estimators with the reference's signatures (policy/base.py:317-440).

    Like the reference class it is a ``torch.nn.Module`` (``state_dict`` / ``train`` / ``eval``
    work, policy/base.py:62-70) and keeps the surrounding ``update`` pipeline
    (``sample -> process_fn -> learn -> post_process_fn``, policy/base.py:288-315).  The
    constructor takes the reference's keyword arguments (policy/base.py:72-109); action
    scaling / bounding is the Collector's concern (``map_action``) and is only recorded here.
    """

    def __init__(
        self,
        observation_space: Any = None,
        action_space: Any = None,
        action_scaling: bool = False,
        action_bound_method: Optional[str] = None,
        lr_scheduler: Optional[Any] = None,
    ) -> None:
        if action_bound_method is not None:
            assert action_bound_method in ("clip", "tanh")
        super().__init__()
        if isinstance(action_space, list):  # vector envs (policy/base.py:91-92)
            action_space = action_space[0]
        self.observation_space = observation_space
        self.action_space = action_space
        self.agent_id = 0
        self.updating = False
        self.action_scaling = action_scaling
        self.action_bound_method = action_bound_method
        self.lr_scheduler = lr_scheduler

    def set_agent_id(self, agent_id: int) -> None:
        self.agent_id = agent_id

    def __getstate__(self) -> Dict[str, Any]:
        """Captured CUDA graphs hold device addresses: they are rebuilt after unpickling."""
        state = dict(self.__dict__)
        if state.get("_graphs") is not None:
            state["_graphs"] = None
        if state.get("_graphed"):
            state["_graphed"] = {}
        return state

    # ---- data-parallel learners --------------------------------------------------
    def set_data_parallel(self, enabled: bool = True, group: Any = None) -> None:
        """One process per GPU, each with its own env streams / samples: after ``backward`` the
        gradients are summed over ``group`` and divided by its size, so every rank applies the
        same step.  Per-minibatch advantage normalisation stays rank-local (ppo.py:109-113
        normalises over the minibatch it sees)."""
        self._dp_group = (group,) if enabled else None

    def _sync_grads(self, params: Any) -> None:
        dp = getattr(self, "_dp_group", None)
        if dp is not None:
            allreduce_mean_grads(params, dp[0])

    def forward(self, batch: Batch, state: Any = None, **kwargs: Any) -> Batch:
        raise NotImplementedError

    # ---- update pipeline (policy/base.py:244-315) -----------------------------------------------
    def process_fn(self, batch: Batch, buffer: Any, indices: ArrayLike) -> Batch:
        return batch

    def learn(self, batch: Batch, **kwargs: Any) -> Dict[str, Any]:
        raise NotImplementedError

    def post_process_fn(self, batch: Batch, buffer: Any, indices: ArrayLike) -> None:
        if hasattr(buffer, "update_weight") and hasattr(batch, "weight"):
            buffer.update_weight(indices, batch.weight)  # policy/base.py:285-286

    def update(self, sample_size: int, buffer: Optional[Any], **kwargs: Any) -> Dict[str, Any]:
        if buffer is None:
            return {}
        batch, indices = buffer.sample(sample_size)
        self.updating = True
        batch = self.process_fn(batch, buffer, indices)
        result = self.learn(batch, **kwargs)
        self.post_process_fn(batch, buffer, indices)
        if self.lr_scheduler is not None:
            self.lr_scheduler.step()
        self.updating = False
        return result

    # ---- value mask (policy/base.py:317-335) ------------------------------------------------------
    @staticmethod
    def value_mask(buffer: Any, indices: ArrayLike):
        """``~buffer.terminated[indices]``: whether obs_next of buffer[indices] is valid."""
        term = buffer.terminated
        if isinstance(term, torch.Tensor):
            idx = indices if isinstance(indices, torch.Tensor) else torch.as_tensor(
                np.asarray(indices), device=term.device)
            out = ~term[idx.to(term.device)]
            return out if isinstance(indices, torch.Tensor) else out.cpu().numpy()
        return ~term[indices]

    # ---- GAE (policy/base.py:337-384) ---------------------------------------------------------------
    @staticmethod
    def compute_episodic_return(
        batch: Batch,
        buffer: Any,
        indices: ArrayLike,
        v_s_: Optional[ArrayLike] = None,
        v_s: Optional[ArrayLike] = None,
        gamma: float = 0.99,
        gae_lambda: float = 0.95,
        out_dtype: Optional[torch.dtype] = None,
    ) -> Tuple[ArrayLike, ArrayLike]:
        """GAE advantages and TD(lambda) returns of ``batch == buffer[indices]``.

        Same contract as the reference: episodes are delimited by the batch's
        ``terminated | truncated`` plus ``buffer.unfinished_index()``; ``v_s_`` is masked by
        ``~buffer.terminated[indices]``; ``v_s=None`` means ``roll(v_s_, 1)``; ``v_s_=None``
        requires ``gae_lambda == 1`` and means zeros.  Returns ``(returns, advantage)``.

        ``out_dtype`` defaults to float64 for NumPy callers (what numba returns) and to the
        dtype of ``v_s`` for tensor callers (what ``to_torch_as`` would produce, a2c.py:115).
        """
        dev = _device_of(buffer)
        tensor_io = isinstance(v_s_, torch.Tensor) or isinstance(v_s, torch.Tensor) or (
            v_s_ is None and v_s is None and isinstance(batch.rew, torch.Tensor))
        canon = buffer.is_canonical(indices) if hasattr(buffer, "is_canonical") else None
        if (not tensor_io and canon is not None and v_s_ is not None and v_s is not None
                and canon[0] * canon[1] >= HOST_CALL_MIN_N):
            out = _episodic_return_host_call(batch, dev, canon, v_s_, v_s, gamma, gae_lambda,
                                             out_dtype)
            if out is not None:
                return out
        rew = _float_dev(batch.rew, dev)
        n = rew.numel()
        off, last, lens, B, sub, buf_done, buf_term = _buffer_layout(buffer, dev)
        idx = _dev(indices, dev, torch.int64)
        if idx.ndim != 1:
            idx = idx.reshape(-1)
        assert idx.numel() == n, "indices must address one buffer slot per batch element"

        if v_s_ is None:
            assert np.isclose(gae_lambda, 1.0)  # policy/base.py:372
            vn = torch.zeros_like(rew)  # :373
            masked_vn = vn
        else:
            vn = _float_dev(v_s_, dev)
            masked_vn = None  # the kernel applies the value mask (:376)
        if v_s is None:  # :377 -- roll of the MASKED v_s_
            if masked_vn is None:
                masked_vn = vn * (buf_term[idx] == 0).to(vn.dtype)
            vs = torch.roll(masked_vn, 1)
        else:
            vs = _float_dev(v_s, dev)
        if vs.dtype != vn.dtype:  # one value dtype per launch: widen, never narrow
            vs, vn = vs.to(torch.float64), vn.to(torch.float64)
        assert vs.numel() == n and vn.numel() == n

        if out_dtype is None:
            out_dtype = vs.dtype if tensor_io else torch.float64
        adv = torch.empty(n, dtype=out_dtype, device=dev)
        ret = torch.empty(n, dtype=out_dtype, device=dev)
        lib = L.lib()
        if n == 0:
            pass
        elif canon is not None and v_s_ is not None:
            # canonical [E, T] rollout (buffer.sample(0) of an un-wrapped VectorReplayBuffer):
            # batch order IS row-major [E, T], every row ends on a done/unfinished step, so
            # the rows kernel runs without indices, unfinished list or look-back.  The batch's
            # own rew/flags are read (a caller may have rewritten batch.rew, e.g. GAIL).
            E, T = canon
            term = L.as_u8(_flat(_dev(batch.terminated, dev, torch.bool)))
            trunc = L.as_u8(_flat(_dev(batch.truncated, dev, torch.bool)))
            L.check(lib.ts_gae_rows(
                L.ptr(rew), L.ptr(vs), L.ptr(vn), L.ptr(term), L.ptr(trunc), E, T, T, T,
                float(gamma), float(gae_lambda), L.dtype_code(rew), L.dtype_code(vs),
                L.TS_F64 if out_dtype == torch.float64 else L.TS_F32, L.ptr(adv), L.ptr(ret),
                L.stream()), "ts_gae_rows")
        elif (v_s_ is not None and n >= LISTING_MIN_N and hasattr(buffer, "listing_runs")
              and buffer.listing_runs(indices) is not None):
            # The batch is the buffer's own chronological listing (sample(0) of ragged or
            # wrapped sub-buffers): runs of consecutive slots.  The unfinished tails become B
            # cut flags at batch positions (ts_mark_unfinished) and the value mask is the
            # batch's own `terminated`, as on the canonical path, so the flat kernel runs
            # WITHOUT the N indices: 21 instead of 30 bytes per transition.
            runs = buffer.listing_runs(indices)
            term = L.as_u8(_flat(_dev(batch.terminated, dev, torch.bool)))
            cuts = L.as_u8(_flat(_dev(batch.truncated, dev, torch.bool))).clone()
            L.check(lib.ts_mark_unfinished(
                L.ptr(buf_done), L.ptr(last), L.ptr(lens), L.ptr(off), L.ptr(runs["start"]),
                L.ptr(runs["first"]), B, n, L.ptr(cuts), L.stream()), "ts_mark_unfinished")
            ws = _workspace(lib.ts_gae_flat_workspace_bytes(n), dev)
            L.check(lib.ts_gae_flat(
                L.ptr(rew), L.ptr(vs), L.ptr(vn), L.ptr(term), L.ptr(cuts), None, None, None,
                L.ptr(off), B, sub, n, float(gamma), float(gae_lambda), L.dtype_code(rew),
                L.dtype_code(vs), L.TS_F64 if out_dtype == torch.float64 else L.TS_F32,
                L.ptr(adv), L.ptr(ret), L.ptr(ws), ws.numel(), L.stream()), "ts_gae_flat")
        else:
            term = L.as_u8(_flat(_dev(batch.terminated, dev, torch.bool)))
            trunc = L.as_u8(_flat(_dev(batch.truncated, dev, torch.bool)))
            unfinished = torch.empty(B, dtype=torch.int64, device=dev)
            L.check(lib.ts_unfinished_index(L.ptr(buf_done), L.ptr(last), L.ptr(lens), B,
                                            L.ptr(unfinished), L.stream()),
                    "ts_unfinished_index")
            nbytes = lib.ts_gae_flat_workspace_bytes(n)
            ws = _workspace(nbytes, dev)
            L.check(lib.ts_gae_flat(
                L.ptr(rew), L.ptr(vs), L.ptr(vn), L.ptr(term), L.ptr(trunc), L.ptr(idx),
                L.ptr(buf_term) if v_s_ is not None else None, L.ptr(unfinished), L.ptr(off), B,
                sub, n, float(gamma), float(gae_lambda), L.dtype_code(rew), L.dtype_code(vs),
                L.TS_F64 if out_dtype == torch.float64 else L.TS_F32, L.ptr(adv), L.ptr(ret),
                L.ptr(ws), ws.numel(), L.stream()), "ts_gae_flat")
        if tensor_io:
            return ret, adv
        return ret.cpu().numpy(), adv.cpu().numpy()

    # ---- n-step (policy/base.py:386-440) -------------------------------------------------------------
    @staticmethod
    def compute_nstep_return(
        batch: Batch,
        buffer: Any,
        indice: ArrayLike,
        target_q_fn: Callable[[Any, ArrayLike], torch.Tensor],
        gamma: float = 0.99,
        n_step: int = 1,
        rew_norm: bool = False,
    ) -> Batch:
        r"""n-step TD target, stored in ``batch.returns`` as a tensor shaped ``(bsz, -1)`` with
        the dtype/device of ``target_q_fn``'s output (policy/base.py:431, :437).

        ``target_q_fn(buffer, terminal)`` receives ``terminal`` as the same array kind as
        ``indice`` (CUDA tensor in -> CUDA tensor, NumPy in -> NumPy).
        """
        assert not rew_norm, \
            "Reward normalization in computing n-step returns is unsupported now."
        dev = _device_of(buffer)
        off, last, lens, B, sub, buf_done, buf_term = _buffer_layout(buffer, dev)
        rew = buffer.rew if isinstance(buffer, ReplayBuffer) else _dev(buffer.rew, dev)
        if rew.dtype not in (torch.float32, torch.float64):
            rew = rew.to(torch.float64)
        idx = _flat(_dev(indice, dev, torch.int64))
        bsz = idx.numel()
        sh = L.shim()
        if sh is not None and isinstance(buffer, ReplayBuffer) and bsz > 0:
            # the same two C-ABI calls from C++ (csrc/shim.cpp)
            st = buffer._state()
            terminal, fl = sh.nstep_chain(idx, rew.contiguous(), buf_done, buf_term, st["packed"],
                                          B, sub, float(gamma), int(n_step))
            terminal_arg = terminal if isinstance(indice, torch.Tensor) else terminal.cpu().numpy()
            with torch.no_grad():
                target_q_torch = target_q_fn(buffer, terminal_arg)  # (bsz, ?)  :430
            tq_dev = target_q_torch.reshape(bsz, -1).to(dev)
            if tq_dev.dtype not in (torch.float32, torch.float64):
                tq_dev = tq_dev.to(torch.float32)
            out = sh.nstep_combine(tq_dev.contiguous(), fl)
            batch.returns = out.to(device=target_q_torch.device, dtype=target_q_torch.dtype)
            if hasattr(batch, "weight"):  # prio buffer update  :438-439
                w = batch.weight
                if not isinstance(w, torch.Tensor):
                    w = torch.from_numpy(np.asarray(w))
                batch.weight = w.to(device=target_q_torch.device, dtype=target_q_torch.dtype)
            return batch
        terminal = torch.empty(bsz, dtype=torch.int64, device=dev)
        ret_sum = torch.empty(bsz, dtype=torch.float64, device=dev)
        coef = torch.empty(bsz, dtype=torch.float64, device=dev)
        lib = L.lib()
        L.check(lib.ts_nstep_chain(
            L.ptr(idx), bsz, L.ptr(rew.contiguous()), L.dtype_code(rew), L.ptr(buf_done),
            L.ptr(buf_term), L.ptr(off), B, sub, L.ptr(last), L.ptr(lens), float(gamma),
            int(n_step), L.ptr(terminal), L.ptr(ret_sum), L.ptr(coef), L.stream()),
            "ts_nstep_chain")
        terminal_arg = terminal if isinstance(indice, torch.Tensor) else terminal.cpu().numpy()
        with torch.no_grad():
            target_q_torch = target_q_fn(buffer, terminal_arg)  # (bsz, ?)  :430
        tq = target_q_torch.reshape(bsz, -1)
        tq_dev = tq.to(dev)
        if tq_dev.dtype not in (torch.float32, torch.float64):
            tq_dev = tq_dev.to(torch.float32)
        tq_dev = tq_dev.contiguous()
        out = torch.empty_like(tq_dev)
        L.check(lib.ts_nstep_combine(L.ptr(tq_dev), L.dtype_code(tq_dev), bsz, tq_dev.shape[1],
                                     L.ptr(ret_sum), L.ptr(coef), L.ptr(out), L.stream()),
                "ts_nstep_combine")
        batch.returns = out.to(device=target_q_torch.device, dtype=target_q_torch.dtype)  # :437
        if hasattr(batch, "weight"):  # prio buffer update  :438-439
            w = batch.weight
            if not isinstance(w, torch.Tensor):
                w = torch.from_numpy(np.asarray(w))
            batch.weight = w.to(device=target_q_torch.device, dtype=target_q_torch.dtype)
        return batch
