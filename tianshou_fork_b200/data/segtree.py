"""Device-resident ``SegmentTree`` -- drop-in for ``tianshou.data.SegmentTree``
(tianshou/data/utils/segtree.py:7-137).

Same constructor, ``len``, ``tree[index]``, ``tree[index] = value``, ``reduce(start, end)`` and
``get_prefix_sum_idx(value)``; the f64 heap lives in HBM and the three numba kernels are
replaced by the CUDA kernels in ``csrc/segtree.cu``.  Inputs may be Python scalars, lists,
NumPy arrays or torch tensors; NumPy in -> NumPy out, CUDA tensor in -> CUDA tensor out (no
host sync on the tensor path).  Reference behaviours kept on purpose:

* duplicate indices in ``__setitem__``: the last occurrence wins (segtree.py:40-49);
* ``get_prefix_sum_idx`` mutates its array argument in place (:134) and returns a Python
  int for a scalar (:80-85);
* range contracts (:53, :79) raise ``AssertionError``; ``tree[size_beyond_heap]`` raises
  ``IndexError`` (test/base/test_buffer.py:504-505).

On the NumPy path the asserts are evaluated on the host before launch like the reference;
on the tensor path they are evaluated on the device into an error word that is checked by
``check_errors()`` (or automatically when ``eager_check`` is set).
"""
from __future__ import annotations

from typing import Optional, Union

import numpy as np
import torch

from .. import _lib as L

Index = Union[int, list, np.ndarray, torch.Tensor]


class SegmentTree:
    def __init__(self, size: int, device: Union[str, torch.device, None] = None,
                 eager_check: bool = False) -> None:
        L.lib()  # fail loudly without the extension or a GPU
        bound = 1
        while bound < size:  # segtree.py:22-24
            bound *= 2
        self._size = int(size)
        self._bound = bound
        self.device = torch.device(device if device is not None else
                                   f"cuda:{torch.cuda.current_device()}")
        self._heap = torch.zeros(bound * 2, dtype=torch.float64, device=self.device)
        # Prioritized buffers update this tree lazily (ts_per_update_lazy): the levels above the
        # update kernel's range roots may be stale until the next sampling launch rebuilds them.
        # Every other reader goes through the `_value` property, which finishes the top first.
        self._top_stale = False
        # scratch of the update kernel (see include/ts_b200.h): all-zero between calls
        self._stamp = torch.zeros(max(size, 1), dtype=torch.int32, device=self.device)
        self._counter = torch.zeros(4, dtype=torch.int32, device=self.device)
        self._err = torch.zeros(1, dtype=torch.int32, device=self.device)
        self._scalar_out = torch.zeros(1, dtype=torch.float64, device=self.device)
        self.eager_check = eager_check

    def __len__(self) -> int:
        return self._size

    @property
    def _value(self) -> torch.Tensor:
        """The f64 heap of ``2 * bound`` nodes (segtree.py:27), complete."""
        if self._top_stale:
            self._top_stale = False
            L.check(L.lib().ts_segtree_finish_top(self._heap.data_ptr(), self._bound,
                                                  L.stream()), "ts_segtree_finish_top")
        return self._heap

    # ---- helpers ---------------------------------------------------------------------------
    def _dev_index(self, index: Index) -> torch.Tensor:
        if isinstance(index, torch.Tensor):
            t = index.to(device=self.device, dtype=torch.int64)
        else:
            t = torch.as_tensor(np.asarray(index, dtype=np.int64), device=self.device)
        return t.reshape(-1).contiguous()

    def check_errors(self) -> None:
        """Raise the AssertionError the reference would have raised (one D2H sync)."""
        flags = int(self._err.item())
        if flags:
            self._err.zero_()
            if flags & L.TS_FLAG_INDEX_RANGE:
                raise AssertionError("SegmentTree index out of range [0, size)")  # :53
            if flags & L.TS_FLAG_PEER_TIMEOUT:
                raise RuntimeError("sharded PER: a peer's record did not arrive in time")
            if flags & L.TS_FLAG_PREFIX_RANGE:
                raise AssertionError("prefix-sum value outside [0, tree.reduce())")  # :79

    # ---- reads -----------------------------------------------------------------------------
    def __getitem__(self, index: Index):
        """``self._value[index + bound]`` (segtree.py:33-35)."""
        if isinstance(index, torch.Tensor):
            return self._value[index.to(self.device) + self._bound]
        idx = np.asarray(index) + self._bound
        if np.any(idx >= 2 * self._bound) or np.any(idx < -2 * self._bound):
            raise IndexError(f"index {index} is out of bounds for SegmentTree of bound "
                             f"{self._bound}")
        out = self._value[torch.as_tensor(idx, device=self.device)].cpu().numpy()
        return out.item() if np.ndim(index) == 0 else out

    def reduce(self, start: int = 0, end: Optional[int] = None) -> float:
        if start == 0 and end is None:  # :58-59
            return float(self._value[1].item())
        if end is None:
            end = self._size
        if end < 0:
            end += self._size
        L.check(L.lib().ts_segtree_reduce(L.ptr(self._value), self._bound, int(start), int(end),
                                          L.ptr(self._scalar_out), L.stream()),
                "ts_segtree_reduce")
        return float(self._scalar_out.item())

    def root(self) -> torch.Tensor:
        """``reduce()`` as a 0-d device tensor (no host sync)."""
        return self._value[1]

    # ---- writes ----------------------------------------------------------------------------
    def __setitem__(self, index: Index, value) -> None:
        host_path = not isinstance(index, torch.Tensor)
        if host_path:
            idx_np = np.asarray(index)
            assert np.all(0 <= idx_np) and np.all(idx_np < self._size)  # :53
        idx = self._dev_index(index)
        m = idx.numel()
        if m == 0:
            raise IndexError("index 0 is out of bounds for axis 0 with size 0")  # index[0], :102
        if isinstance(value, torch.Tensor):
            val = value.to(self.device)
        else:
            val = torch.as_tensor(np.asarray(value), device=self.device)
        if val.dtype not in (torch.float32, torch.float64):
            val = val.to(torch.float64)
        if val.numel() == 1:
            val, stride = val.reshape(1).contiguous(), 0
        else:
            assert val.numel() == m, "value must be a scalar or match index"
            val, stride = val.reshape(-1).contiguous(), 1
        L.check(L.lib().ts_segtree_update(
            L.ptr(self._value), self._size, self._bound, L.ptr(idx), L.ptr(val),
            L.dtype_code(val), stride, m, L.ptr(self._stamp), L.ptr(self._counter),
            L.ptr(self._err), L.stream()), "ts_segtree_update")
        if self.eager_check and not host_path:
            self.check_errors()

    # ---- queries ---------------------------------------------------------------------------
    def get_prefix_sum_idx(self, value):
        """Minimum index i with ``value <= sum(arr[:i+1])`` (segtree.py:66-85)."""
        if isinstance(value, torch.Tensor):
            v = value
            assert v.is_cuda and v.is_contiguous() and v.dtype in (torch.float32, torch.float64)
            out = torch.empty(v.shape, dtype=torch.int64, device=self.device)
            L.check(L.lib().ts_segtree_prefix_idx(
                L.ptr(self._value), self._bound, L.ptr(v), L.dtype_code(v), v.numel(), 0,
                L.ptr(out), L.ptr(self._err), L.stream()), "ts_segtree_prefix_idx")
            if self.eager_check:
                self.check_errors()
            return out
        total = self.reduce()
        assert np.all(np.asarray(value) >= 0.0) and np.all(np.asarray(value) < total)  # :79
        single = not isinstance(value, np.ndarray)
        arr = np.array([value], dtype=np.float64) if single else value
        if arr.dtype not in (np.float32, np.float64):
            raise TypeError("get_prefix_sum_idx expects float32/float64 values")
        v = torch.as_tensor(np.ascontiguousarray(arr), device=self.device)
        out = torch.empty(v.shape, dtype=torch.int64, device=self.device)
        L.check(L.lib().ts_segtree_prefix_idx(
            L.ptr(self._value), self._bound, L.ptr(v), L.dtype_code(v), v.numel(), 0,
            L.ptr(out), L.ptr(self._err), L.stream()), "ts_segtree_prefix_idx")
        idx = out.cpu().numpy()
        if not single:
            arr[...] = v.cpu().numpy()  # the reference mutates the caller's array (:134)
        return idx.item() if single else idx

    # ---- pickling: host arrays, like the reference object -----------------------------------
    def __getstate__(self):
        return {"_size": self._size, "_bound": self._bound,
                "_value": self._value.cpu().numpy(), "device": str(self.device),
                "eager_check": self.eager_check}

    def __setstate__(self, state):
        self.__init__(state["_size"], device=state["device"], eager_check=state["eager_check"])
        self._value.copy_(torch.from_numpy(state["_value"]))
