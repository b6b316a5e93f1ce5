"""``ReplayBufferManager`` / ``VectorReplayBuffer`` with device-resident storage
(tianshou/data/buffer/manager.py:10-192, vecbuf.py:15-37).

N sub-buffers (one per env stream) share one contiguous storage; sub-buffer ``k`` owns the
global index range ``[_offset[k], _offset[k+1])`` and holds a VIEW of the storage
(manager.py:60-62).  An episode never crosses a sub-buffer (manager.py:269-276), which is
what lets env streams shard across GPUs with no communication.

Differences from the reference that are not visible through the API:
* ``add`` advances all rings with vectorised NumPy instead of a Python loop over
  sub-buffers (manager.py:136-145), same results;
* ``prev`` / ``next`` are O(len(index)) kernels instead of O(buffer_num * len(index)).
"""
from __future__ import annotations

from typing import Any, Dict, List, Optional, Tuple, Union

import numpy as np
import torch

from .. import _lib as L
from .batch import Batch, create_value
from .buffer import ReplayBuffer, _alloc_missing, _host


class ReplayBufferManager(ReplayBuffer):
    """Contiguous storage shared by a list of equal-option ReplayBuffers."""

    def __init__(self, buffer_list: List[ReplayBuffer]) -> None:
        self.buffer_num = len(buffer_list)
        self.buffers = np.array(buffer_list, dtype=object)
        offset, size = [], 0
        buffer_type = type(self.buffers[0])
        kwargs = self.buffers[0].options
        for buf in self.buffers:
            assert buf._meta.is_empty()
            assert isinstance(buf, buffer_type) and buf.options == kwargs
            offset.append(size)
            size += buf.maxsize
        self._offset = np.array(offset, dtype=np.int64)
        self._extend_offset = np.array(offset + [size], dtype=np.int64)
        self._lengths = np.zeros_like(self._offset)
        self._sub_sizes = np.array([b.maxsize for b in self.buffers], dtype=np.int64)
        # equal sub-buffers (VectorReplayBuffer) -> kernels find the sub-buffer by division
        self._equal_sub = int(self._sub_sizes[0]) if np.all(
            self._sub_sizes == self._sub_sizes[0]) else 0
        # ring table [B, 2] = (index, size); children alias their row
        self._rings = np.zeros((self.buffer_num, 2), dtype=np.int64)
        self._ep_rew = np.zeros(self.buffer_num, dtype=np.float64)
        self._ep_len = np.zeros(self.buffer_num, dtype=np.int64)
        self._ep_idx = np.zeros(self.buffer_num, dtype=np.int64)
        self._child_last = np.zeros((self.buffer_num, 1), dtype=np.int64)  # local last_index
        for k, buf in enumerate(self.buffers):
            buf._ring = self._rings[k]
            buf.last_index = self._child_last[k]
            buf._parent = self
        first = self.buffers[0]
        super().__init__(size=size, device=first.device, rew_dtype=first.rew_dtype, **kwargs)

    def __len__(self) -> int:
        return int(self._lengths.sum())

    def reset(self, keep_statistics: bool = False) -> None:
        self.last_index = self._offset.copy()  # manager.py:55
        self._lengths = np.zeros_like(self._offset)
        self._rings[:] = 0
        if not keep_statistics:
            self._ep_rew[:] = 0.0
            self._ep_len[:] = 0
            self._ep_idx[:] = 0
        self._child_last[:] = 0
        self._version += 1

    def _set_batch_for_children(self) -> None:
        for offset, buf in zip(self._offset, self.buffers):
            buf.set_batch(self._meta[int(offset):int(offset) + buf.maxsize])  # views

    def set_batch(self, batch: Batch) -> None:
        super().set_batch(batch)
        self._set_batch_for_children()

    def _layout(self):
        return (self._extend_offset, np.asarray(self.last_index, dtype=np.int64),
                np.asarray(self._lengths, dtype=np.int64), self._equal_sub)

    def update(self, buffer: ReplayBuffer) -> np.ndarray:
        """The ReplayBufferManager cannot be updated by any buffer (manager.py:100-102)."""
        raise NotImplementedError

    # ---- add ---------------------------------------------------------------------------------
    def _advance_rings(self, ids: np.ndarray, rew: np.ndarray, done: np.ndarray):
        """Vectorised ``_add_index`` over the sub-buffers in ``ids`` (base.py:195-214 applied
        per sub-buffer as manager.py:136-145 does)."""
        sub = self._sub_sizes[ids]
        ptr_local = self._rings[ids, 0].copy()
        self._rings[ids, 1] = np.minimum(self._rings[ids, 1] + 1, sub)
        self._rings[ids, 0] = (ptr_local + 1) % sub
        ptrs = ptr_local + self._offset[ids]
        self._ep_rew[ids] += rew
        self._ep_len[ids] += 1
        ep_rews = np.where(done, self._ep_rew[ids], 0.0)
        ep_lens = np.where(done, self._ep_len[ids], 0)
        ep_idxs = self._ep_idx[ids] + self._offset[ids]
        fin = ids[done]
        self._ep_rew[fin] = 0.0
        self._ep_len[fin] = 0
        self._ep_idx[fin] = self._rings[fin, 0]
        self.last_index[ids] = ptrs  # manager.py:144
        self._lengths[ids] = self._rings[ids, 1]  # :145
        self._child_last[ids, 0] = ptr_local  # children keep a local last_index (base.py:202)
        return ptrs, ep_rews, ep_lens, ep_idxs

    def add(self, batch: Batch, buffer_ids: Optional[Union[np.ndarray, List[int]]] = None
            ) -> Tuple[np.ndarray, np.ndarray, np.ndarray, np.ndarray]:
        """Add one transition per listed sub-buffer (manager.py:104-161)."""
        batch = self._prepare_add(batch, self._reserved_keys, stacked=True)
        if buffer_ids is None:
            buffer_ids = np.arange(self.buffer_num)
        ids = np.asarray(buffer_ids, dtype=np.int64)
        rew, done = _host(batch.rew), np.asarray(batch.done, dtype=bool)
        if rew.ndim == 1 and len(np.unique(ids)) == len(ids):
            ptrs, ep_rews, ep_lens, ep_idxs = self._advance_rings(ids, rew, done)
        else:  # vector rewards or repeated ids: sequential, exactly the reference loop
            ptrs, ep_lens, ep_rews, ep_idxs = [], [], [], []
            for batch_idx, buffer_id in enumerate(ids):
                buf = self.buffers[buffer_id]
                ptr, ep_rew, ep_len, ep_idx = buf._add_index(rew[batch_idx], done[batch_idx])
                ptrs.append(ptr + self._offset[buffer_id])
                ep_lens.append(ep_len)
                ep_rews.append(ep_rew)
                ep_idxs.append(ep_idx + self._offset[buffer_id])
                self.last_index[buffer_id] = ptr + self._offset[buffer_id]
                self._lengths[buffer_id] = len(buf)
            ptrs, ep_rews = np.array(ptrs), np.array(ep_rews)
            ep_lens, ep_idxs = np.array(ep_lens), np.array(ep_idxs)
        self._version += 1
        batch = self._widen(batch)
        if self._meta.is_empty():
            self._meta = create_value(batch, self.maxsize, stack=False, device=self.device)
            self._set_batch_for_children()
        elif not set(batch.keys()).issubset(self._meta.keys()):
            _alloc_missing(self._meta, batch, self.maxsize, False, self.device)
            self._set_batch_for_children()
        self._meta[ptrs] = batch
        return ptrs, ep_rews, ep_lens, ep_idxs

    # ---- sampling ------------------------------------------------------------------------------
    def unfinished_index(self) -> np.ndarray:
        """manager.py:68-74 (concatenated per-sub-buffer unfinished indices)."""
        return super().unfinished_index()

    def canonical_rows(self) -> Optional[Tuple[int, int]]:
        """(E, T) when every sub-buffer holds the same number T of steps stored from its
        column 0 in chronological order (un-wrapped): the layout ``ts_gae_rows`` consumes."""
        if not self._equal_sub:
            return None
        sizes, idx = self._rings[:, 1], self._rings[:, 0]
        T = int(sizes[0])
        if T == 0 or np.any(sizes != T):
            return None
        if not np.all((idx == T) | ((idx == 0) & (T == self._equal_sub))):
            return None
        return self.buffer_num, T

    def _all_indices_device(self) -> torch.Tensor:
        """``sample_indices(0)``: every sub-buffer's ``[_index.._size) ++ [0.._index)`` plus its
        offset, concatenated (manager.py:187-192 over base.py:285-289), built on the device."""
        dev = self.device
        canon = self.canonical_rows()
        if canon is not None:
            E, T = canon
            off = torch.arange(E, device=dev, dtype=torch.int64) * self._equal_sub
            out = (off[:, None] + torch.arange(T, device=dev, dtype=torch.int64)[None]).reshape(-1)
            self._canon = {"ptr": out.data_ptr(), "n": out.numel(), "E": E, "T": T,
                           "version": self._version}
            return out
        sizes = torch.from_numpy(self._rings[:, 1].copy()).to(dev)
        start = torch.from_numpy(self._rings[:, 0].copy()).to(dev)
        off = torch.from_numpy(self._offset).to(dev)
        total = int(self._rings[:, 1].sum())
        k = torch.repeat_interleave(torch.arange(self.buffer_num, device=dev), sizes,
                                    output_size=total)
        first = torch.cumsum(sizes, 0) - sizes
        j = torch.arange(total, device=dev) - first[k]
        return (start[k] + j) % torch.clamp(sizes[k], min=1) + off[k]

    def is_canonical(self, indices: Any) -> Optional[Tuple[int, int]]:
        """Whether ``indices`` is the tensor the last ``sample_indices(0)`` returned for the
        current contents in canonical layout; gives (E, T)."""
        c = getattr(self, "_canon", None)
        if (c is not None and isinstance(indices, torch.Tensor)
                and c["version"] == self._version and indices.data_ptr() == c["ptr"]
                and indices.numel() == c["n"]):
            return c["E"], c["T"]
        return None

    def sample_indices(self, batch_size: int) -> torch.Tensor:
        """manager.py:163-192; random draws come from the host ``np.random`` stream in the
        reference's call order so a seeded run draws identical indices."""
        if batch_size < 0:
            return torch.empty(0, dtype=torch.int64, device=self.device)
        if self._sample_avail and self.stack_num > 1:
            all_indices = torch.cat([
                buf.sample_indices(0) + int(offset)
                for offset, buf in zip(self._offset, self.buffers)])
            if batch_size == 0:
                return all_indices
            pick = np.random.choice(all_indices.numel(), batch_size)
            return all_indices[torch.from_numpy(pick).to(self.device)]
        if batch_size == 0:
            return self._all_indices_device()
        buffer_idx = np.random.choice(
            self.buffer_num, batch_size, p=self._lengths / self._lengths.sum())
        sample_num = np.bincount(buffer_idx, minlength=self.buffer_num)
        parts = []
        for k in np.flatnonzero(sample_num):  # sub-buffers with zero draws make no RNG call
            parts.append(np.random.choice(int(self._rings[k, 1]), int(sample_num[k]))
                         + self._offset[k])
        idx = np.concatenate(parts) if parts else np.array([], int)
        return torch.from_numpy(idx.astype(np.int64)).to(self.device)


class VectorReplayBuffer(ReplayBufferManager):
    """``buffer_num`` equal ReplayBuffers of ``ceil(total_size / buffer_num)`` slots
    (tianshou/data/buffer/vecbuf.py:15-37)."""

    def __init__(self, total_size: int, buffer_num: int, **kwargs: Any) -> None:
        assert buffer_num > 0
        size = int(np.ceil(total_size / buffer_num))
        buffer_list = [ReplayBuffer(size, **kwargs) for _ in range(buffer_num)]
        super().__init__(buffer_list)
