"""CPU oracle for the return / replay hot path -- TEST INFRASTRUCTURE ONLY.

Thin ctypes wrapper over ``oracle/ts_oracle.c`` (the C restatement of the reference's
five numba kernels) plus numpy restatements of the small amount of host logic around
them.  Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this module; the product package
``tianshou_fork_b200`` never does and has no CPU fallback.

Parity status: PINNED -- ``tests/test_oracle.py`` checks every function here against the
reference's own known-answer vectors and against ``tests/golden/*.npz`` produced by the
imported reference (``tests/golden/make_golden.py``; numba 0.65.0, numpy 2.3.5).

Paths in citations are relative to the reference checkout.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from typing import Optional, Tuple

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libts_oracle.so")
_lib = None

_c_f64p = ctypes.POINTER(ctypes.c_double)
_c_f32p = ctypes.POINTER(ctypes.c_float)
_c_i64p = ctypes.POINTER(ctypes.c_int64)
_c_u8p = ctypes.POINTER(ctypes.c_uint8)


def build(force: bool = False) -> str:
    """Compile ``ts_oracle.c`` with gcc (``make -C oracle``)."""
    src = os.path.join(_HERE, "ts_oracle.c")
    if (
        force
        or not os.path.exists(_LIB_PATH)
        or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src)
    ):
        subprocess.check_call(["make", "-s", "-C", _HERE, "-B"])
    return _LIB_PATH


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(_LIB_PATH)
        _lib.orc_segtree_reduce.restype = ctypes.c_double
        _lib.orc_max_threads.restype = ctypes.c_int
    return _lib


def _p(a: np.ndarray, ty):
    return a.ctypes.data_as(ty)


def _f64(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.float64)


def _u8(a) -> np.ndarray:
    return np.ascontiguousarray(np.asarray(a).astype(np.uint8))


def _i64(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.int64)


def max_threads() -> int:
    return int(lib().orc_max_threads())


# ---------------------------------------------------------------------------------------
# GAE  (tianshou/policy/base.py:337-384, 453-497)
# ---------------------------------------------------------------------------------------
def gae_return(v_s, v_s_, rew, end_flag, gamma: float, gae_lambda: float) -> np.ndarray:
    """``_gae_return`` (policy/base.py:453-497): advantages, f64."""
    rew = _f64(rew)
    n = rew.shape[0]
    out = np.zeros(n, dtype=np.float64)
    end_flag = _u8(end_flag)
    v_s = np.ascontiguousarray(v_s)
    v_s_ = np.ascontiguousarray(v_s_)
    if v_s.dtype == np.float32 and v_s_.dtype == np.float32:
        lib().orc_gae_return_f32(
            _p(v_s, _c_f32p), _p(v_s_, _c_f32p), _p(rew, _c_f64p), _p(end_flag, _c_u8p),
            ctypes.c_double(gamma), ctypes.c_double(gae_lambda), ctypes.c_int64(n),
            _p(out, _c_f64p),
        )
    else:
        v_s, v_s_ = _f64(v_s), _f64(v_s_)
        lib().orc_gae_return(
            _p(v_s, _c_f64p), _p(v_s_, _c_f64p), _p(rew, _c_f64p), _p(end_flag, _c_u8p),
            ctypes.c_double(gamma), ctypes.c_double(gae_lambda), ctypes.c_int64(n),
            _p(out, _c_f64p),
        )
    return out


def compute_episodic_return(
    rew, terminated, truncated, buffer_terminated_at_indices, indices, unfinished_index,
    v_s_=None, v_s=None, gamma: float = 0.99, gae_lambda: float = 0.95,
) -> Tuple[np.ndarray, np.ndarray]:
    """``BasePolicy.compute_episodic_return`` (policy/base.py:337-384).

    ``rew/terminated/truncated`` are the batch's fields, ``buffer_terminated_at_indices``
    is ``buffer.terminated[indices]`` (value_mask, :335) and ``unfinished_index`` is
    ``buffer.unfinished_index()`` (:380).
    """
    rew = np.asarray(rew)
    if v_s_ is None:
        assert np.isclose(gae_lambda, 1.0)  # :372
        v_s_ = np.zeros_like(rew)  # :373
    else:
        v_s_ = np.asarray(v_s_).flatten()
        v_s_ = v_s_ * ~np.asarray(buffer_terminated_at_indices, dtype=bool)  # :376
    v_s = np.roll(v_s_, 1) if v_s is None else np.asarray(v_s).flatten()  # :377
    end_flag = np.logical_or(terminated, truncated)  # :379
    end_flag[np.isin(indices, unfinished_index)] = True  # :380
    adv = gae_return(v_s, v_s_, rew, end_flag, gamma, gae_lambda)  # :381
    return adv + v_s, adv  # :382


def episodic_return_rows(rew, v_s, v_s_, terminated, truncated, gamma, gae_lambda,
                         threads: int = 1) -> Tuple[np.ndarray, np.ndarray]:
    """``compute_episodic_return`` over an ``[E, T]`` rollout of independent streams
    (see ``orc_episodic_return_rows``).  f32 in, f64 out."""
    rew = np.ascontiguousarray(rew, dtype=np.float32)
    v_s = np.ascontiguousarray(v_s, dtype=np.float32)
    v_s_ = np.ascontiguousarray(v_s_, dtype=np.float32)
    terminated, truncated = _u8(terminated), _u8(truncated)
    E, T = rew.shape
    ret = np.empty((E, T), dtype=np.float64)
    adv = np.empty((E, T), dtype=np.float64)
    lib().orc_episodic_return_rows(
        _p(rew, _c_f32p), _p(v_s, _c_f32p), _p(v_s_, _c_f32p), _p(terminated, _c_u8p),
        _p(truncated, _c_u8p), ctypes.c_int64(E), ctypes.c_int64(T),
        ctypes.c_double(gamma), ctypes.c_double(gae_lambda), _p(ret, _c_f64p),
        _p(adv, _c_f64p), ctypes.c_int(threads),
    )
    return ret, adv


# ---------------------------------------------------------------------------------------
# n-step  (tianshou/policy/base.py:386-440, 500-524)
# ---------------------------------------------------------------------------------------
def nstep_return(rew, end_flag, target_q, indices, gamma: float, n_step: int) -> np.ndarray:
    """``_nstep_return`` (policy/base.py:500-524). ``indices`` is ``[n_step, bsz]``."""
    rew, end_flag = _f64(rew), _u8(end_flag)
    target_q = _f64(target_q)
    shape = target_q.shape
    bsz = shape[0]
    tq = target_q.reshape(bsz, -1)
    indices = _i64(indices)
    out = np.zeros_like(tq)
    lib().orc_nstep_return(
        _p(rew, _c_f64p), _p(end_flag, _c_u8p), _p(tq, _c_f64p), _p(indices, _c_i64p),
        ctypes.c_double(gamma), ctypes.c_int64(n_step), ctypes.c_int64(bsz),
        ctypes.c_int64(tq.shape[1]), _p(out, _c_f64p),
    )
    return out.reshape(shape)


# ---------------------------------------------------------------------------------------
# Index arithmetic  (tianshou/data/buffer/base.py:148-168, manager.py:68-98, 259-297)
# ---------------------------------------------------------------------------------------
def prev_index(index, offset, done, last_index, lengths) -> np.ndarray:
    """``_prev_index`` (manager.py:259-277). ``offset`` is ``_extend_offset`` (B+1)."""
    index, offset = _i64(np.atleast_1d(index)), _i64(offset)
    done, last_index, lengths = _u8(done), _i64(last_index), _i64(lengths)
    out = np.zeros_like(index)
    lib().orc_prev_index(
        _p(index, _c_i64p), ctypes.c_int64(index.size), _p(offset, _c_i64p),
        ctypes.c_int64(offset.size - 1), _p(done, _c_u8p), _p(last_index, _c_i64p),
        _p(lengths, _c_i64p), _p(out, _c_i64p),
    )
    return out


def next_index(index, offset, done, last_index, lengths) -> np.ndarray:
    """``_next_index`` (manager.py:280-297)."""
    index, offset = _i64(np.atleast_1d(index)), _i64(offset)
    done, last_index, lengths = _u8(done), _i64(last_index), _i64(lengths)
    out = np.zeros_like(index)
    lib().orc_next_index(
        _p(index, _c_i64p), ctypes.c_int64(index.size), _p(offset, _c_i64p),
        ctypes.c_int64(offset.size - 1), _p(done, _c_u8p), _p(last_index, _c_i64p),
        _p(lengths, _c_i64p), _p(out, _c_i64p),
    )
    return out


def prev_single(index, done, last_index: int, size: int) -> np.ndarray:
    """``ReplayBuffer.prev`` (base.py:153-160): modulo by the CURRENT size."""
    index, done = _i64(np.atleast_1d(index)), _u8(done)
    out = np.zeros_like(index)
    lib().orc_prev_single(_p(index, _c_i64p), ctypes.c_int64(index.size), _p(done, _c_u8p),
                          ctypes.c_int64(last_index), ctypes.c_int64(size), _p(out, _c_i64p))
    return out


def next_single(index, done, last_index: int, size: int) -> np.ndarray:
    """``ReplayBuffer.next`` (base.py:162-168)."""
    index, done = _i64(np.atleast_1d(index)), _u8(done)
    out = np.zeros_like(index)
    lib().orc_next_single(_p(index, _c_i64p), ctypes.c_int64(index.size), _p(done, _c_u8p),
                          ctypes.c_int64(last_index), ctypes.c_int64(size), _p(out, _c_i64p))
    return out


def stack_gather(frames, index, stack_num: int, offset, done, last_index, lengths,
                 from_next: bool = False) -> np.ndarray:
    """``ReplayBuffer.get(index, key, stack_num)`` (tianshou/data/buffer/base.py:348-354):
    ``stack = [val[indices]] + stack; indices = self.prev(indices)`` repeated ``stack_num``
    times, stacked on ``axis = indices.ndim`` (oldest frame first).  ``from_next`` starts
    from ``next(index)`` (the obs_next of ignore_obs_next buffers, base.py:378)."""
    frames = np.asarray(frames)
    idx = _i64(np.asarray(index).reshape(-1))
    if from_next:
        idx = next_index(idx, offset, done, last_index, lengths)
    stack = []
    for _ in range(stack_num):
        stack = [frames[idx]] + stack
        idx = prev_index(idx, offset, done, last_index, lengths)
    return np.stack(stack, axis=1)


class RingState:
    """Bookkeeping of N equal rings inside one contiguous storage.

    Restates the state machine of ``ReplayBuffer._add_index`` (base.py:195-214) and
    ``ReplayBufferManager.add/reset/unfinished_index/sample_indices(0)``
    (manager.py:54-58, 68-74, 104-161, 177-192) for the fields the hot path reads:
    ``done``, ``last_index``, ``_lengths``, per-ring ``_index`` / ``_size``.
    """

    def __init__(self, total_size: int, buffer_num: int) -> None:
        self.buffer_num = buffer_num
        self.sub = int(np.ceil(total_size / buffer_num))  # vecbuf.py:35
        self.maxsize = self.sub * buffer_num
        self.offset = np.arange(buffer_num, dtype=np.int64) * self.sub
        self.extend_offset = np.arange(buffer_num + 1, dtype=np.int64) * self.sub
        self.done = np.zeros(self.maxsize, dtype=bool)
        self.terminated = np.zeros(self.maxsize, dtype=bool)
        self.truncated = np.zeros(self.maxsize, dtype=bool)
        self.rew = np.zeros(self.maxsize, dtype=np.float64)
        self.reset()

    def reset(self) -> None:
        self.last_index = self.offset.copy()  # manager.py:55
        self.lengths = np.zeros(self.buffer_num, dtype=np.int64)  # :56
        self.index = np.zeros(self.buffer_num, dtype=np.int64)  # base.py:137
        self.size = np.zeros(self.buffer_num, dtype=np.int64)

    def add(self, rew, terminated, truncated, buffer_ids=None) -> np.ndarray:
        """Returns ``ptrs`` (manager.py:136-146)."""
        if buffer_ids is None:
            buffer_ids = np.arange(self.buffer_num)
        ptrs = []
        for j, k in enumerate(buffer_ids):
            ptr = int(self.index[k])  # base.py:202
            self.size[k] = min(self.size[k] + 1, self.sub)  # :203
            self.index[k] = (self.index[k] + 1) % self.sub  # :204
            g = ptr + int(self.offset[k])
            self.last_index[k] = g  # manager.py:144
            self.lengths[k] = self.size[k]  # :145
            self.rew[g] = rew[j]
            self.terminated[g] = bool(terminated[j])
            self.truncated[g] = bool(truncated[j])
            self.done[g] = bool(terminated[j]) or bool(truncated[j])  # :123
            ptrs.append(g)
        return np.array(ptrs, dtype=np.int64)

    def unfinished_index(self) -> np.ndarray:
        out = []
        for k in range(self.buffer_num):  # manager.py:68-74 over base.py:148-151
            if self.size[k]:
                last = (self.index[k] - 1) % self.size[k]
                if not self.done[last + self.offset[k]]:
                    out.append(last + self.offset[k])
        return np.array(out, dtype=np.int64)

    def sample_indices_all(self) -> np.ndarray:
        """``sample_indices(0)`` (manager.py:177-192 over base.py:285-289)."""
        parts = []
        for k in range(self.buffer_num):
            i, s = int(self.index[k]), int(self.size[k])
            parts.append(np.concatenate([np.arange(i, s), np.arange(i)]) + self.offset[k])
        return np.concatenate(parts).astype(np.int64)

    def prev(self, index) -> np.ndarray:
        return prev_index(index, self.extend_offset, self.done, self.last_index, self.lengths)

    def next(self, index) -> np.ndarray:
        return next_index(index, self.extend_offset, self.done, self.last_index, self.lengths)


def compute_nstep_return(state: "RingState", indice, target_q_fn, gamma: float,
                         n_step: int) -> np.ndarray:
    """``BasePolicy.compute_nstep_return`` (policy/base.py:386-440) on a ``RingState``.

    ``target_q_fn(terminal) -> [bsz, ...]`` numpy.  Returns f64 ``[bsz, D]``: the
    reference reshapes the target to ``(bsz, -1)`` BEFORE the kernel (:431), so a flat
    ``[bsz]`` target comes back as ``[bsz, 1]``.
    """
    indices = [np.asarray(indice, dtype=np.int64)]
    for _ in range(n_step - 1):  # :423-424
        indices.append(state.next(indices[-1]))
    indices = np.stack(indices)
    terminal = indices[-1]  # :428
    tq = np.asarray(target_q_fn(terminal))
    bsz = len(indice)
    tq = tq.reshape(bsz, -1) * (~state.terminated[terminal]).reshape(-1, 1)  # :431-432
    end_flag = state.done.copy()  # :433
    end_flag[state.unfinished_index()] = True  # :434
    return nstep_return(state.rew, end_flag, tq, indices, gamma, n_step)


# ---------------------------------------------------------------------------------------
# Segment tree  (tianshou/data/utils/segtree.py)
# ---------------------------------------------------------------------------------------
class SegTree:
    """``SegmentTree`` (segtree.py:7-85): f64 heap of ``2 * bound`` nodes."""

    def __init__(self, size: int) -> None:
        bound = 1
        while bound < size:  # :22-24
            bound *= 2
        self._size, self._bound = size, bound
        self._value = np.zeros(bound * 2, dtype=np.float64)

    def __len__(self) -> int:
        return self._size

    def __getitem__(self, index):
        return self._value[index + self._bound]  # :35 (numpy raises IndexError beyond)

    def __setitem__(self, index, value) -> None:
        if isinstance(index, (int, np.integer)):  # :51-52
            index, value = np.array([index]), np.array([value])
        index = np.asarray(index)
        assert np.all(0 <= index) and np.all(index < self._size)  # :53
        idx = _i64(index + self._bound).copy()
        val = np.asarray(value)
        if val.ndim == 0:
            val = np.full(idx.shape, val)
        # f32 values are cast on store into the f64 tree (:91-92, :101)
        val = _f64(val)
        lib().orc_segtree_setitem(_p(self._value, _c_f64p), _p(idx, _c_i64p),
                                  _p(val, _c_f64p), ctypes.c_int64(idx.size))

    def reduce(self, start: int = 0, end: Optional[int] = None) -> float:
        if start == 0 and end is None:  # :58-59
            return float(self._value[1])
        if end is None:
            end = self._size
        if end < 0:
            end += self._size
        return float(lib().orc_segtree_reduce(
            _p(self._value, _c_f64p), ctypes.c_int64(start + self._bound - 1),
            ctypes.c_int64(end + self._bound)))

    def get_prefix_sum_idx(self, value):
        assert np.all(np.asarray(value) >= 0.0) and np.all(np.asarray(value) < self._value[1])
        single = False
        if not isinstance(value, np.ndarray):  # :81-83
            value = np.array([value], dtype=np.float64)
            single = True
        assert value.flags.c_contiguous
        out = np.zeros(value.shape, dtype=np.int64)
        if value.dtype == np.float32:
            lib().orc_segtree_prefix_idx_f32(
                _p(value, _c_f32p), ctypes.c_int64(value.size), ctypes.c_int64(self._bound),
                _p(self._value, _c_f64p), _p(out, _c_i64p))
        else:
            assert value.dtype == np.float64
            lib().orc_segtree_prefix_idx(
                _p(value, _c_f64p), ctypes.c_int64(value.size), ctypes.c_int64(self._bound),
                _p(self._value, _c_f64p), _p(out, _c_i64p))
        return out.item() if single else out


class PER:
    """Priority state of ``PrioritizedReplayBuffer`` (prio.py:22-102) without storage."""

    def __init__(self, size: int, alpha: float, beta: float, weight_norm: bool = True) -> None:
        assert alpha > 0.0 and beta >= 0.0  # :33
        self._alpha, self._beta = alpha, beta
        self._max_prio = self._min_prio = 1.0  # :35
        self.weight = SegTree(size)
        self._eps = np.finfo(np.float32).eps.item()  # :38
        self._weight_norm = weight_norm

    def init_weight(self, index) -> None:
        self.weight[index] = self._max_prio ** self._alpha  # :43

    def sample_indices(self, u: np.ndarray) -> np.ndarray:
        """prio.py:61-62 with the uniform draws ``u`` supplied by the caller."""
        scalar = u * self.weight.reduce()
        return self.weight.get_prefix_sum_idx(scalar)

    def get_weight(self, index):
        return (self.weight[index] / self._min_prio) ** (-self._beta)  # :76

    def batch_weight(self, index):
        w = self.get_weight(index)
        return w / np.max(w) if self._weight_norm else w  # :101

    def update_weight(self, index, new_weight) -> None:
        weight = np.abs(np.asarray(new_weight)) + self._eps  # :86 (stays f32 for f32 input)
        self.weight[index] = weight ** self._alpha  # :87
        self._max_prio = max(self._max_prio, weight.max())  # :88
        self._min_prio = min(self._min_prio, weight.min())  # :89


def per_step_c(tree: SegTree, u, td, alpha, beta, weight_norm, max_prio, min_prio):
    """One fused sample + weight + update step in C (``orc_per_step``), for timing."""
    bsz = u.size
    idx = np.empty(bsz, dtype=np.int64)
    w = np.empty(bsz, dtype=np.float64)
    sv = np.empty(bsz, dtype=np.float64)
    si = np.empty(bsz, dtype=np.int64)
    mx, mn = ctypes.c_double(max_prio), ctypes.c_double(min_prio)
    lib().orc_per_step(
        _p(tree._value, _c_f64p), ctypes.c_int64(tree._bound), _p(_f64(u), _c_f64p),
        _p(np.ascontiguousarray(td, dtype=np.float32), _c_f32p), ctypes.c_int64(bsz),
        ctypes.c_double(alpha), ctypes.c_double(beta), ctypes.c_int(int(weight_norm)),
        ctypes.byref(mx), ctypes.byref(mn), _p(idx, _c_i64p), _p(w, _c_f64p),
        _p(sv, _c_f64p), _p(si, _c_i64p))
    return idx, w, mx.value, mn.value


# ---------------------------------------------------------------------------------------
# PPO loss arithmetic  (tianshou/policy/modelfree/ppo.py:109-144), f64 numpy
# ---------------------------------------------------------------------------------------
def ppo_loss(logp, logp_old, adv, returns, v_s, value, eps_clip: float = 0.2,
             dual_clip: Optional[float] = None, value_clip: bool = False,
             norm_adv: bool = True, eps: float = 1e-8):
    """Returns ``(clip_loss, vf_loss, dlogp, dvalue, adv_used)`` evaluated in f64.

    ``dlogp`` / ``dvalue`` are d(clip_loss)/d(logp) and d(vf_loss)/d(value), derived by
    hand from the expressions at ppo.py:114-139 (torch.min/max/clamp subgradients pick
    the branch torch picks: ties in min/max split evenly is NOT modelled -- inputs in the
    tests avoid exact ties).
    """
    logp, logp_old = np.asarray(logp, np.float64), np.asarray(logp_old, np.float64)
    adv, returns = np.asarray(adv, np.float64), np.asarray(returns, np.float64)
    v_s, value = np.asarray(v_s, np.float64), np.asarray(value, np.float64)
    n = adv.size
    if norm_adv:  # :109-113, torch.std is unbiased
        mean, std = adv.mean(), adv.std(ddof=1)
        adv = (adv - mean) / (std + eps)
    ratio = np.exp(logp - logp_old)  # :114-116
    surr1 = ratio * adv  # :118
    clipped = np.clip(ratio, 1.0 - eps_clip, 1.0 + eps_clip)
    surr2 = clipped * adv  # :119-122
    clip1 = np.minimum(surr1, surr2)
    in_range = (ratio >= 1.0 - eps_clip) & (ratio <= 1.0 + eps_clip)
    # d min(surr1, surr2)/d ratio: adv where surr1 is the min (or equal), else adv*in_range
    take1 = surr1 <= surr2
    dclip1 = np.where(take1, adv, adv * in_range)
    if dual_clip:  # :123-126
        clip2 = np.maximum(clip1, dual_clip * adv)
        obj = np.where(adv < 0, clip2, clip1)
        dobj = np.where(adv < 0, np.where(clip1 >= dual_clip * adv, dclip1, 0.0), dclip1)
    else:
        obj, dobj = clip1, dclip1  # :128
    clip_loss = -obj.mean()
    dlogp = -(dobj * ratio) / n
    if value_clip:  # :131-137
        dv = value - v_s
        v_clip = v_s + np.clip(dv, -eps_clip, eps_clip)
        vf1 = (returns - value) ** 2
        vf2 = (returns - v_clip) ** 2
        vf_loss = np.maximum(vf1, vf2).mean()
        inr = (dv >= -eps_clip) & (dv <= eps_clip)
        d1 = -2.0 * (returns - value)
        d2 = -2.0 * (returns - v_clip) * inr
        dvalue = np.where(vf1 >= vf2, d1, d2) / n
    else:
        vf_loss = ((returns - value) ** 2).mean()  # :139
        dvalue = -2.0 * (returns - value) / n
    return clip_loss, vf_loss, dlogp, dvalue, adv


# ---------------------------------------------------------------------------------------------
# Sharded PER: G power-of-two-aligned shard trees == one global tree
# ---------------------------------------------------------------------------------------------
def global_tree_from_shards(shards) -> "SegTree":
    """The reference's single global tree (manager.py:209-214) whose leaves are the shards'
    leaves laid end to end; requires equal power-of-two shard bounds."""
    G, b = len(shards), shards[0]._bound
    assert G & (G - 1) == 0 and all(s._bound == b and s._size == b for s in shards)
    g = SegTree(G * b)
    g[np.arange(G * b)] = np.concatenate([s._value[b:] for s in shards])
    return g


def sharded_route(u, roots):
    """Top of the global tree rebuilt from the shard roots in its pairwise order
    (segtree.py:104), then ``_get_prefix_sum_idx`` (:122-137) over those log2(G) levels.
    Returns (total, shard[M], residual value[M]) for ``scalar = u * total`` (prio.py:61)."""
    roots = np.asarray(roots, dtype=np.float64)
    G = roots.size
    top = np.zeros(2 * G)
    top[G:] = roots
    for i in range(G - 1, 0, -1):
        top[i] = top[2 * i] + top[2 * i + 1]
    total = top[1] if G > 1 else roots[0]
    v = np.asarray(u, dtype=np.float64) * total
    node = np.ones(v.shape, dtype=np.int64)
    while node[0] < G:
        node = node * 2
        lsons = top[node]
        direct = lsons < v
        v = v - lsons * direct
        node = node + direct
    return total, node - G, v
