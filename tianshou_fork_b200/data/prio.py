"""Prioritized replay with a device-resident sum tree -- drop-ins for
``tianshou.data.PrioritizedReplayBuffer`` (tianshou/data/buffer/prio.py:9-105),
``PrioritizedReplayBufferManager`` (manager.py:195-214) and
``PrioritizedVectorReplayBuffer`` (vecbuf.py:40-68).

State that the reference keeps as Python floats (``_max_prio``, ``_min_prio``, prio.py:35,
88-89) lives in a 2-element f64 device tensor so that ``sample -> learn -> update_weight``
never synchronises with the host; the properties of the same names read it back on demand.

Reference quirks reproduced on purpose:
* ``_min_prio`` / ``_max_prio`` track ``|td| + eps`` BEFORE ``** alpha`` while leaves hold the
  value after it (prio.py:86-89 vs :76);
* the vector manager rebuilds itself from ``buffer_list[0].options`` which never carries
  ``weight_norm`` (prio.py:39, manager.py:211-214), so a PrioritizedVectorReplayBuffer ALWAYS
  normalises the importance weights;
* ``PrioritizedVectorReplayBuffer.set_beta`` only reaches the children (vecbuf.py:66-68) while
  ``get_weight`` reads the manager's own ``_beta``: annealing beta has no effect there.

Bit-exactness is defined at the SegmentTree boundary: identical leaves + identical uniform
draws => identical sampled indices.  Leaf values themselves go through ``pow`` (f32 in the
reference when td is f32) and are compared with a tolerance.
"""
from __future__ import annotations

from typing import Any, List, Optional, Sequence, Tuple, Union

import numpy as np
import torch

from .. import _lib as L
from .batch import Batch
from .buffer import IndexLike, ReplayBuffer
from .manager import ReplayBufferManager
from .segtree import SegmentTree


class PrioritizedReplayBuffer(ReplayBuffer):
    """Prioritized Experience Replay, arXiv:1511.05952 (prio.py:9-105)."""

    def __init__(self, size: int, alpha: float, beta: float, weight_norm: bool = True,
                 **kwargs: Any) -> None:
        ReplayBuffer.__init__(self, size, **kwargs)
        assert alpha > 0.0 and beta >= 0.0  # prio.py:33
        self._alpha, self._beta = alpha, beta
        self.weight = SegmentTree(size, device=self.device)
        # {max_prio, min_prio} (prio.py:35) + whether leaves are currently f32-typed powers
        self._prio = torch.ones(2, dtype=torch.float64, device=self.device)
        self._prio_f32 = False
        self._per_scratch = torch.zeros(4, dtype=torch.int32, device=self.device)  # 16 bytes
        self.options.update(alpha=alpha, beta=beta)
        self._weight_norm = weight_norm
        self.device_rng = False  # True: draw the uniforms with torch on the device

    # Python-float views of the device scalars (one D2H sync each; not used on the hot path)
    @property
    def _max_prio(self) -> float:
        return float(self._prio[0].item())

    @property
    def _min_prio(self) -> float:
        return float(self._prio[1].item())

    def init_weight(self, index: IndexLike) -> None:
        """``weight[index] = max_prio ** alpha`` (prio.py:42-43)."""
        idx = self.weight._dev_index(index)
        if idx.numel() == 0:
            return
        self._cached_weight = None  # the tree changes: weights of an earlier draw are stale
        t = self.weight
        L.check(L.lib().ts_per_init_weight(
            L.ptr(t._value), t._size, t._bound, L.ptr(idx), idx.numel(), float(self._alpha),
            L.ptr(self._prio), L.TS_F32 if self._prio_f32 else L.TS_F64, L.ptr(t._stamp),
            L.ptr(t._counter), L.ptr(t._err), L.stream()), "ts_per_init_weight")

    def update(self, buffer: ReplayBuffer) -> np.ndarray:
        indices = super().update(buffer)
        self.init_weight(indices)
        return indices

    def add(self, batch: Batch, buffer_ids: Optional[Union[np.ndarray, List[int]]] = None
            ) -> Tuple[np.ndarray, np.ndarray, np.ndarray, np.ndarray]:
        ptr, ep_rew, ep_len, ep_idx = super().add(batch, buffer_ids)
        self.init_weight(ptr)
        return ptr, ep_rew, ep_len, ep_idx

    def add_device(self, batch: Batch, buffer_ids: Optional[Union[np.ndarray, List[int]]] = None):
        """Device-side ``add`` (manager ``add_device``) followed by ``init_weight`` of the slots
        it wrote, as ``add`` does (prio.py:50-57); the slot indices never leave the device."""
        if not isinstance(self, ReplayBufferManager):
            raise NotImplementedError("add_device needs a buffer manager (VectorReplayBuffer)")
        out = ReplayBufferManager.add_device(self, batch, buffer_ids)
        self.init_weight(out[0])
        return out

    # ---- sampling ------------------------------------------------------------------------------
    def _uniforms(self, batch_size: int) -> torch.Tensor:
        if self.device_rng:
            return torch.rand(batch_size, dtype=torch.float64, device=self.device)
        return torch.from_numpy(np.random.rand(batch_size)).to(self.device)  # prio.py:61

    def sample_indices(self, batch_size: int) -> torch.Tensor:
        """Stratified-by-priority indices (prio.py:59-64).  The importance weights of the
        draw are computed in the same launch and cached for ``__getitem__``."""
        if batch_size > 0 and len(self) > 0:
            return self.sample_with_uniforms(self._uniforms(batch_size))
        return super().sample_indices(batch_size)

    def sample_with_uniforms(self, u: torch.Tensor) -> torch.Tensor:
        """``get_prefix_sum_idx(u * weight.reduce())`` for caller-supplied uniforms."""
        assert u.is_cuda and u.dtype == torch.float64 and u.is_contiguous()
        m = u.numel()
        idx = torch.empty(m, dtype=torch.int64, device=self.device)
        w = torch.empty(m, dtype=torch.float64, device=self.device)
        t = self.weight
        # rebuilds the levels update_weight left stale, in shared memory, and writes them back
        L.check(L.lib().ts_per_sample_lazy(
            t._heap.data_ptr(), t._bound, L.ptr(u), m, L.ptr(self._prio), float(self._beta),
            int(self._weight_norm), L.ptr(idx), L.ptr(w), L.ptr(t._err), L.stream()),
            "ts_per_sample_lazy")
        t._top_stale = False
        self._cached_weight = (idx, w)
        return idx

    # ---- host-array callers (the reference's calling convention: NumPy in, NumPy out) ----------
    def _pinned(self, name: str, nbytes: int) -> torch.Tensor:
        """Reusable pinned staging buffers (one per direction and purpose)."""
        cache = self.__dict__.setdefault("_pin", {})
        t = cache.get(name)
        if t is None or t.numel() < nbytes:
            t = cache[name] = torch.empty(max(nbytes, 1 << 16), dtype=torch.uint8).pin_memory()
        return t

    def sample_indices_numpy(self, batch_size: int) -> Tuple[np.ndarray, np.ndarray]:
        """``sample_indices(batch_size)`` and the importance weights of the draw as NumPy arrays
        (what the reference's ``buffer.sample`` hands a host-side learner: indices + ``weight``,
        prio.py:59-64, 91-102) with ONE upload (the ``np.random.rand`` draws, prio.py:61) and ONE
        download (indices and weights share a device buffer) instead of a copy per array."""
        m = int(batch_size)
        assert m > 0 and len(self) > 0
        stage = self._pinned("u", 8 * m)[: 8 * m].view(torch.float64)
        stage.numpy()[:] = np.random.rand(m)  # prio.py:61
        dev = self.__dict__.setdefault("_host_io", {})
        if dev.get("m") != m:
            dev.clear()
            dev.update(m=m, u=torch.empty(m, dtype=torch.float64, device=self.device),
                       out=torch.empty(2 * m, dtype=torch.int64, device=self.device))
        u, out = dev["u"], dev["out"]
        t = self.weight
        sh = L.shim()
        if sh is not None:  # the same copies and launch issued from C++ (csrc/shim.cpp)
            host = self._pinned("out", 16 * m)[: 16 * m].view(torch.int64)
            sh.per_sample_host(t._heap, t._bound, self._prio, t._err, stage, u, out, host, m,
                               float(self._beta), bool(self._weight_norm))
            t._top_stale = False
            self._cached_weight = None
            h = host.numpy()
            return h[:m].copy(), h[m:].view(np.float64).copy()
        u.copy_(stage, non_blocking=True)
        idx, w = out[:m], out[m:].view(torch.float64)
        L.check(L.lib().ts_per_sample_lazy(
            t._heap.data_ptr(), t._bound, u.data_ptr(), m, self._prio.data_ptr(),
            float(self._beta), int(self._weight_norm), idx.data_ptr(), w.data_ptr(),
            t._err.data_ptr(), L.stream()), "ts_per_sample_lazy")
        t._top_stale = False
        self._cached_weight = None
        host = self._pinned("out", 16 * m)[: 16 * m].view(torch.int64)
        host.copy_(out, non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()
        h = host.numpy()
        return h[:m].copy(), h[m:].view(np.float64).copy()

    def _update_host_shim(self, index: np.ndarray, new_weight: np.ndarray) -> bool:
        """``update_weight`` on host arrays through csrc/shim.cpp: pack into the pinned staging
        buffer, then ONE C++ call (upload + ts_per_update_lazy).  False: shim absent."""
        sh = L.shim()
        if sh is None:
            return False
        idx = np.ascontiguousarray(index, dtype=np.int64).reshape(-1)
        td = np.ascontiguousarray(new_weight).reshape(-1)
        if td.dtype not in (np.float32, np.float64):
            td = td.astype(np.float64)
        m = idx.size
        assert td.size == m
        f32 = td.dtype == np.float32
        nb = 8 * m + td.itemsize * m
        stage = self._pinned("upd", nb)
        stage[: 8 * m].view(torch.int64).numpy()[:] = idx
        stage[8 * m: nb].view(torch.float32 if f32 else torch.float64).numpy()[:] = td
        dev = self.__dict__.setdefault("_host_io_upd", {})
        key = ("upd_dev", nb)
        buf = dev.get(key)
        if buf is None:
            dev.clear()
            buf = dev[key] = torch.empty(nb, dtype=torch.uint8, device=self.device)
        t = self.weight
        heap = t._value  # (a top left stale by a graph replay is finished before it is built on)
        sh.per_update_host(heap, t._size, t._bound, self._prio, t._stamp, t._counter, t._err,
                           stage, buf, m, f32, float(self._alpha))
        ev = self.__dict__.get("_upd_ev")
        if ev is None:
            ev = self.__dict__["_upd_ev"] = torch.cuda.Event()
        ev.record()  # the staging buffer may be rewritten once this copy has run
        self._cached_weight = None
        t._top_stale = True
        self._prio_f32 = f32
        return True

    def _upload_update(self, index: np.ndarray, new_weight: np.ndarray):
        """Host indices + host td-errors -> device, as one pinned buffer and one copy."""
        idx = np.ascontiguousarray(index, dtype=np.int64).reshape(-1)
        td = np.ascontiguousarray(new_weight).reshape(-1)
        if td.dtype not in (np.float32, np.float64):
            td = td.astype(np.float64)
        m = idx.size
        assert td.size == m
        nb = 8 * m + td.itemsize * m
        stage = self._pinned("upd", nb)
        stage[: 8 * m].view(torch.int64).numpy()[:] = idx
        tdt = torch.float32 if td.dtype == np.float32 else torch.float64
        stage[8 * m: nb].view(tdt).numpy()[:] = td
        ev = self.__dict__.get("_upd_ev")
        key = ("upd_dev", nb)
        dev = self.__dict__.setdefault("_host_io_upd", {})
        buf = dev.get(key)
        if buf is None:
            dev.clear()
            buf = dev[key] = torch.empty(nb, dtype=torch.uint8, device=self.device)
        buf.copy_(stage[:nb], non_blocking=True)
        if ev is None:
            ev = self.__dict__["_upd_ev"] = torch.cuda.Event()
        ev.record()  # the staging buffer may be rewritten once this copy has been issued AND run
        return buf[: 8 * m].view(torch.int64), buf[8 * m:].view(tdt), ev

    def get_weight(self, index: IndexLike):
        """Importance weight ``(p_j / p_min) ** (-beta)`` (prio.py:66-76), un-normalised."""
        return self._weights(index, normalise=False)

    def _weights(self, index: IndexLike, normalise: bool):
        idx = self.weight._dev_index(index)
        w = torch.empty(idx.numel(), dtype=torch.float64, device=self.device)
        t = self.weight
        if idx.numel():
            L.check(L.lib().ts_per_weight(
                L.ptr(t._value), t._bound, L.ptr(idx), idx.numel(), L.ptr(self._prio),
                float(self._beta), int(normalise), L.ptr(w), L.ptr(self._per_scratch),
                L.stream()), "ts_per_weight")
        if isinstance(index, torch.Tensor):
            return w.reshape(index.shape)
        out = w.cpu().numpy().reshape(np.shape(index))
        return out.item() if np.ndim(index) == 0 else out

    def update_weight(self, index: IndexLike, new_weight: Union[np.ndarray, torch.Tensor]
                      ) -> None:
        """``weight[index] = (|new_weight| + eps) ** alpha`` plus max/min tracking
        (prio.py:78-89), one launch, no host sync when given CUDA tensors."""
        if (isinstance(index, np.ndarray) and isinstance(new_weight, np.ndarray) and index.size
                and type(self).update_weight is PrioritizedReplayBuffer.update_weight):
            # host arrays in: one packed pinned upload instead of two pageable ones
            ev = self.__dict__.get("_upd_ev")
            if ev is not None:
                ev.synchronize()  # the previous call's copy has left the staging buffer
            if self._update_host_shim(index, new_weight):
                return
            idx, td, _ = self._upload_update(index, new_weight)
        else:
            idx = self.weight._dev_index(index)
            if isinstance(new_weight, torch.Tensor):
                td = new_weight.detach().to(self.device)
            else:
                td = torch.as_tensor(np.asarray(new_weight), device=self.device)
            if td.dtype not in (torch.float32, torch.float64):
                td = td.to(torch.float64)
            td = td.reshape(-1).contiguous()
        assert td.numel() == idx.numel()
        if idx.numel() == 0:
            return
        self._cached_weight = None
        t = self.weight
        # stops at the update kernel's range roots; the next sampling launch (or any reader of
        # `weight._value`) finishes the levels above them
        heap = t._value  # (a top left stale by a graph replay is finished before it is built on)
        L.check(L.lib().ts_per_update_lazy(
            heap.data_ptr(), t._size, t._bound, L.ptr(idx), L.ptr(td), L.dtype_code(td),
            idx.numel(), float(self._alpha), L.ptr(self._prio), L.ptr(t._stamp),
            L.ptr(t._counter), L.ptr(t._err), L.stream()), "ts_per_update_lazy")
        t._top_stale = True
        self._prio_f32 = td.dtype == torch.float32

    def __getitem__(self, index: Union[slice, IndexLike]) -> Batch:
        if isinstance(index, slice):
            indices = self.sample_indices(0) if index == slice(None) \
                else self._as_device_index(np.arange(self.maxsize)[:len(self)][index])
        else:
            indices = self._as_device_index(index)
        batch = super().__getitem__(indices)
        cached = getattr(self, "_cached_weight", None)
        if cached is not None and cached[0] is index:
            batch.weight = cached[1]  # computed by the sampling launch
        else:
            batch.weight = self._weights(indices, normalise=self._weight_norm)  # prio.py:100-101
        self._cached_weight = None
        return batch

    def set_beta(self, beta: float) -> None:
        self._beta = beta

    def check_errors(self) -> None:
        self.weight.check_errors()


class PrioritizedReplayBufferManager(PrioritizedReplayBuffer, ReplayBufferManager):
    """One global tree over the whole capacity (manager.py:195-214)."""

    def __init__(self, buffer_list: Sequence[PrioritizedReplayBuffer]) -> None:
        ReplayBufferManager.__init__(self, buffer_list)  # type: ignore
        kwargs = dict(buffer_list[0].options)
        first = buffer_list[0]
        for buf in buffer_list:
            del buf.weight  # manager.py:212-213
        # weight_norm is not part of `options` -> default True (reference quirk, see module doc)
        meta, rings_version = self._meta, self._version
        PrioritizedReplayBuffer.__init__(self, self.maxsize, device=first.device,
                                         rew_dtype=first.rew_dtype, **kwargs)
        self._meta = meta
        self._version = rings_version + 1


class PrioritizedVectorReplayBuffer(PrioritizedReplayBufferManager):
    """``buffer_num`` equal PrioritizedReplayBuffers (vecbuf.py:40-68)."""

    def __init__(self, total_size: int, buffer_num: int, **kwargs: Any) -> None:
        assert buffer_num > 0
        size = int(np.ceil(total_size / buffer_num))
        buffer_list = [PrioritizedReplayBuffer(size, **kwargs) for _ in range(buffer_num)]
        super().__init__(buffer_list)

    def set_beta(self, beta: float) -> None:
        for buffer in self.buffers:  # vecbuf.py:66-68: the manager's own _beta is untouched
            buffer.set_beta(beta)
