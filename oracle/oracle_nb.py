"""TEST / BENCH INFRASTRUCTURE ONLY -- the reference's hot path restated with the reference's own
technology (numba ``@njit`` array kernels + NumPy glue), so that the CPU baseline timed on the
GPU box's host cores costs what the reference itself costs there.

``oracle/ts_oracle.c`` is the parity checker: a scalar C restatement, which is a *harder* CPU
baseline than the reference (no NumPy temporaries, no Python).  ``/root/reference`` cannot travel
to the GPU box, so ``bench.py``'s ``cpu_baseline`` needs a stand-in that spends time where the
reference spends it: whole-array NumPy expressions around five jitted loops.  Each function
below cites the reference lines it follows; ``scripts/time_reference.py`` times it next to the
imported reference in the build container (``profiles/r02_reference_cpu.json``) and
``tests/test_oracle.py`` checks it value-for-value against the C oracle / the goldens.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU legs may import this module.
"""
from __future__ import annotations

from typing import Optional, Tuple

import numpy as np
from numba import njit


# ---- tianshou/policy/base.py:453-497 ----------------------------------------------------------
@njit
def _gae(v_s, v_s_, rew, end_flag, gamma, gae_lambda):
    out = np.zeros(rew.shape)
    delta = rew + v_s_ * gamma - v_s                       # :491
    discount = (1.0 - end_flag) * (gamma * gae_lambda)     # :492
    acc = 0.0
    for i in range(len(rew) - 1, -1, -1):                  # :494-496
        acc = delta[i] + discount[i] * acc
        out[i] = acc
    return out


def compute_episodic_return(rew, terminated, truncated, buf_terminated, indices, unfinished,
                            v_s_=None, v_s=None, gamma=0.99, gae_lambda=0.95
                            ) -> Tuple[np.ndarray, np.ndarray]:
    """policy/base.py:337-384 with the buffer reduced to ``buf_terminated`` (value_mask, :335) and
    ``unfinished`` (``buffer.unfinished_index()``, :380)."""
    if v_s_ is None:
        assert np.isclose(gae_lambda, 1.0)                 # :372
        v_s_ = np.zeros_like(rew)                          # :373
    else:
        v_s_ = v_s_.flatten()                              # :375
        v_s_ = v_s_ * ~buf_terminated[indices]             # :376 / :335
    v_s = np.roll(v_s_, 1) if v_s is None else v_s.flatten()  # :377
    end_flag = np.logical_or(terminated, truncated)        # :379
    end_flag[np.isin(indices, unfinished)] = True          # :380
    advantage = _gae(v_s, v_s_, rew, end_flag, gamma, gae_lambda)  # :381
    return advantage + v_s, advantage                      # :382-384


# ---- tianshou/policy/base.py:500-524 ----------------------------------------------------------
@njit
def _nstep(rew, end_flag, target_q, indices, gamma, n_step):
    gamma_buffer = np.ones(n_step + 1)
    for i in range(1, n_step + 1):
        gamma_buffer[i] = gamma_buffer[i - 1] * gamma      # :509-510
    shape = target_q.shape
    bsz = shape[0]
    target_q = target_q.reshape(bsz, -1)                   # :514
    returns = np.zeros(target_q.shape)
    gammas = np.full(indices[0].shape, n_step)
    for n in range(n_step - 1, -1, -1):                    # :517-522
        now = indices[n]
        gammas[end_flag[now] > 0] = n + 1
        returns[end_flag[now] > 0] = 0.0
        returns = rew[now].reshape(bsz, 1) + gamma * returns
    target_q = target_q * gamma_buffer[gammas].reshape(bsz, 1) + returns  # :523
    return target_q.reshape(shape)


# ---- tianshou/data/buffer/manager.py:259-297 --------------------------------------------------
@njit
def _next_index(index, offset, done, last_index, lengths):
    index = index % offset[-1]                             # :288
    out = np.zeros_like(index)
    for start, end, cur_len, last in zip(offset[:-1], offset[1:], lengths, last_index):
        mask = (start <= index) & (index < end)            # :290
        cur_len = max(1, cur_len)                          # :292
        if np.sum(mask) > 0:
            sub = index[mask]
            end_flag = done[sub] | (sub == last)           # :295
            out[mask] = (sub - start + 1 - end_flag) % cur_len + start  # :296
    return out


def compute_nstep_return(rew, done, terminated, offset, last_index, lengths, unfinished,
                         indice, target_q_fn, gamma, n_step) -> np.ndarray:
    """policy/base.py:386-440 over the manager's bookkeeping arrays."""
    indices = [indice]
    for _ in range(n_step - 1):                            # :423-424
        indices.append(_next_index(indices[-1], offset, done, last_index, lengths))
    indices = np.stack(indices)
    terminal = indices[-1]                                 # :428
    tq = target_q_fn(terminal).reshape(len(indice), -1)    # :430-431
    tq = tq * (~terminated[terminal]).reshape(-1, 1)       # :432
    end_flag = done.copy()                                 # :433
    end_flag[unfinished] = True                            # :434
    return _nstep(rew, end_flag, tq, indices, gamma, n_step)


# ---- tianshou/data/utils/segtree.py:87-137 ----------------------------------------------------
@njit
def _setitem(tree, index, value):
    tree[index] = value                                    # :101
    while index[0] > 1:                                    # :102-104
        index //= 2
        tree[index] = tree[index * 2] + tree[index * 2 + 1]


@njit
def _get_prefix_sum_idx(value, bound, sums):
    index = np.ones(value.shape, dtype=np.int64)           # :128
    while index[0] < bound:                                # :129-135
        index *= 2
        lsons = sums[index]
        direct = lsons < value
        value -= lsons * direct
        index += direct
    index -= bound
    return index


class PER:
    """The priority side of ``PrioritizedReplayBuffer`` (data/buffer/prio.py:22-105) over a
    ``SegmentTree`` (data/utils/segtree.py:21-85): the work of one sample + update step."""

    def __init__(self, size: int, alpha: float, beta: float, weight_norm: bool = True) -> None:
        bound = 1
        while bound < size:                                # segtree.py:22-24
            bound *= 2
        self._size, self._bound = size, bound
        self._value = np.zeros([bound * 2])
        self._alpha, self._beta = alpha, beta
        self._max_prio = self._min_prio = 1.0              # prio.py:35
        self._eps = np.finfo(np.float32).eps.item()        # :38
        self._weight_norm = weight_norm

    def set_leaves(self, index: np.ndarray, value: np.ndarray) -> None:
        assert np.all(0 <= index) and np.all(index < self._size)  # segtree.py:53
        _setitem(self._value, index + self._bound, value)  # :54

    def sample_indices(self, batch_size: int, u: Optional[np.ndarray] = None) -> np.ndarray:
        if u is None:
            u = np.random.rand(batch_size)
        scalar = u * self._value[1]                        # prio.py:61, segtree.py:58-59
        assert np.all(scalar >= 0.0) and np.all(scalar < self._value[1])  # segtree.py:79
        return _get_prefix_sum_idx(scalar, self._bound, self._value)      # :84

    def batch_weight(self, index: np.ndarray) -> np.ndarray:
        weight = (self._value[index + self._bound] / self._min_prio) ** (-self._beta)  # prio.py:76
        return weight / np.max(weight) if self._weight_norm else weight              # :101

    def update_weight(self, index: np.ndarray, new_weight: np.ndarray) -> None:
        weight = np.abs(new_weight) + self._eps            # prio.py:86
        self.set_leaves(index, weight ** self._alpha)      # :87
        self._max_prio = max(self._max_prio, weight.max())  # :88
        self._min_prio = min(self._min_prio, weight.min())  # :89


def warm() -> None:
    """Compile every kernel for the dtypes the bench uses (the reference does the same in
    ``BasePolicy._compile`` / ``SegmentTree._compile``)."""
    f64 = np.array([0.0, 1.0])
    f32 = f64.astype(np.float32)
    b = np.array([False, True])
    _gae(f64, f64, f64, b, 0.1, 0.1)
    _gae(f32, f32, f64, b, 0.1, 0.1)
    _gae(f32, f32, f32, b, 0.1, 0.1)
    i64 = np.array([[0, 1]], dtype=np.int64)
    _nstep(f64, b, f32.reshape(-1, 1), i64, 0.1, 1)
    _nstep(f64, b, f64.reshape(-1, 1), i64, 0.1, 1)
    _next_index(i64[0].copy(), np.array([0, 2]), b, np.array([1]), np.array([2]))
    _setitem(np.zeros(4), np.array([2, 3]), f64)
    _setitem(np.zeros(4), np.array([2, 3]), f32)
    _get_prefix_sum_idx(f64.copy() * 0, 2, np.ones(4))
