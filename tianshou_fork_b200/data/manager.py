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

import ctypes
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

    def __setstate__(self, state: Dict[str, Any]) -> None:
        """After unpickling, the sub-buffers alias the manager's bookkeeping rows and hold views
        of its storage again (manager.py:60-62); NumPy views do not survive a pickle."""
        super().__setstate__(state)
        for k, buf in enumerate(self.buffers):
            buf.__dict__["_ring"] = self._rings[k]
            buf.__dict__["last_index"] = self._child_last[k]
            buf.__dict__["_parent"] = self
        if not self._meta.is_empty():
            self._set_batch_for_children()

    def reset(self, keep_statistics: bool = False) -> None:
        self.last_index = self._offset.copy()  # manager.py:55
        self._lengths = np.zeros_like(self._offset)
        self._rings[:] = 0
        if not keep_statistics:
            self._ep_rew[:] = 0.0
            self._ep_len[:] = 0
            self._ep_idx[:] = 0
        self._child_last[:] = 0
        self._dev_add = None
        self._ep_on_device = False
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
        if getattr(self, "_ep_on_device", False):
            self._pull_episode_stats()
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
        if not self._store_packed(ptrs, batch):
            self._meta[ptrs] = batch
        return ptrs, ep_rews, ep_lens, ep_idxs

    # ---- device-side add ------------------------------------------------------
    def _pull_episode_stats(self) -> None:
        """Running episode statistics live on the device while ``add_device`` is in use."""
        st = getattr(self, "_dev_add", None)
        if st is not None and getattr(self, "_ep_on_device", False):
            self._ep_rew[:] = st["ep_rew"].cpu().numpy()
            self._ep_len[:] = st["ep_len"].cpu().numpy()
            self._ep_idx[:] = st["ep_idx"].cpu().numpy()
        self._ep_on_device = False

    def _device_add_state(self) -> Dict[str, torch.Tensor]:
        st = getattr(self, "_dev_add", None)
        if st is None or self._dev_add_version != self._version:
            self._pull_episode_stats()  # no-op unless the device copy is the newer one
            dev = self.device
            up = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
            st = {"offset": up(self._extend_offset), "ring": up(self._rings),
                  "last_index": up(np.asarray(self.last_index, dtype=np.int64)),
                  "lengths": up(self._lengths), "ep_rew": up(self._ep_rew),
                  "ep_len": up(self._ep_len), "ep_idx": up(self._ep_idx),
                  "err": torch.zeros(1, dtype=torch.int32, device=dev)}
            st["ptrs"] = [st[k].data_ptr() for k in ("offset", "ring", "last_index", "lengths",
                                                     "ep_rew", "ep_len", "ep_idx", "err")]
            self._dev_add = st
        return st

    def add_device(self, batch: Batch, buffer_ids: Optional[Union[np.ndarray, List[int]]] = None
                   ) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor, torch.Tensor]:
        """``add`` for steps produced on the GPU: every value of ``batch`` is a CUDA tensor with
        leading dimension ``len(buffer_ids)``; ONE launch advances the rings, the episode
        statistics and stores every key (``ts_add_step``), with no host read of ``rew`` /
        ``done``.  Same bookkeeping as manager.py:104-161 over base.py:195-214; returns
        ``(ptrs, ep_rew, ep_len, ep_idx)`` as DEVICE tensors.  ``buffer_ids`` (host) must not
        repeat a sub-buffer.  The ring pointers are mirrored on the host (they do not depend on
        the data), so ``sample_indices`` / ``prev`` / ``next`` / ``__len__`` work as usual."""
        import ctypes
        src = batch.__dict__
        assert all(k in src for k in ("obs", "act", "rew", "terminated", "truncated"))
        vals: Dict[str, torch.Tensor] = {}
        for key in self._reserved_keys:
            v = src.get(key)
            if v is None or (isinstance(v, Batch) and v.is_empty()):
                continue
            if not isinstance(v, torch.Tensor):
                raise TypeError(f"add_device stores tensor-valued keys only (got {key})")
            vals[key] = v
        vals["done"] = torch.logical_or(vals["terminated"], vals["truncated"])
        if self._save_only_last_obs:
            vals["obs"] = vals["obs"][:, -1]
        if not self._save_obs_next:
            vals.pop("obs_next", None)
        elif self._save_only_last_obs and "obs_next" in vals:
            vals["obs_next"] = vals["obs_next"][:, -1]
        full = buffer_ids is None
        if full:
            M = self.buffer_num
        else:
            ids = np.asarray(buffer_ids, dtype=np.int64)
            M = int(ids.size)
            assert len(np.unique(ids)) == M, "add_device: a sub-buffer may appear once per call"
        rew = vals["rew"]
        assert rew.ndim == 1 and rew.shape[0] == M
        if rew.dtype != self.rew_dtype:
            vals["rew"] = rew.to(self.rew_dtype)  # stored as float (manager.py:150)
        if self._meta.is_empty() or not all(k in self._meta.__dict__ for k in vals):
            sample = Batch(**{k: (v.to(torch.bool) if k in ("done", "terminated", "truncated")
                                  else v) for k, v in vals.items()})
            if self._meta.is_empty():
                self._meta = create_value(sample, self.maxsize, stack=False, device=self.device)
            else:
                _alloc_missing(self._meta, sample, self.maxsize, False, self.device)
            self._set_batch_for_children()
            self._dev_add_keys = None
        st = self._device_add_state()
        names = tuple(sorted(vals))
        cache = getattr(self, "_dev_add_keys", None)
        if cache is None or cache["names"] != names:
            stores = [self._meta[k] for k in names]
            cache = {"names": names, "stores": stores, "n": len(names),
                     "dst": (ctypes.c_void_p * len(names))(*[
                         L.ptr(L.as_u8(t)) for t in stores]),
                     "rb": (ctypes.c_int64 * len(names))(*[
                         int(np.prod(t.shape[1:], dtype=np.int64)) * t.element_size()
                         for t in stores])}
            self._dev_add_keys = cache
        keep, srcs = [], []
        for k, store in zip(names, cache["stores"]):
            v = vals[k]
            if v.dtype != store.dtype:
                v = v.to(store.dtype)
            if not v.is_contiguous():
                v = v.contiguous()
            assert v.shape[0] == M and v.shape[1:] == store.shape[1:] and v.is_cuda, k
            keep.append(v)
            srcs.append(v.data_ptr())
        c_src = (ctypes.c_void_p * cache["n"])(*srcs)
        dev = self.device
        g_ids = None if full else torch.from_numpy(ids).to(dev)
        rew = keep[names.index("rew")]
        done = keep[names.index("done")]
        out_i = torch.empty((3, M), dtype=torch.int64, device=dev)
        ep_rew_out = torch.empty(M, dtype=torch.float64, device=dev)
        L.check(L.lib().ts_add_step(
            cache["n"], c_src, cache["dst"], cache["rb"], M,
            g_ids.data_ptr() if g_ids is not None else None, rew.data_ptr(), L.dtype_code(rew),
            done.data_ptr(), st["ptrs"][0], self.buffer_num, st["ptrs"][1], st["ptrs"][2],
            st["ptrs"][3], st["ptrs"][4], st["ptrs"][5], st["ptrs"][6], out_i[0].data_ptr(),
            ep_rew_out.data_ptr(), out_i[1].data_ptr(), out_i[2].data_ptr(), st["ptrs"][7],
            L.stream()), "ts_add_step")
        # host mirror of the pointer bookkeeping (data-independent): _add_index lines :197-199
        r = self._rings
        if full:
            ptr_local = r[:, 0].copy()
            np.minimum(r[:, 1] + 1, self._sub_sizes, out=r[:, 1])
            r[:, 0] = (ptr_local + 1) % self._sub_sizes
            np.add(ptr_local, self._offset, out=self.last_index)
            self._lengths[:] = r[:, 1]
            self._child_last[:, 0] = ptr_local
        else:
            sub = self._sub_sizes[ids]
            ptr_local = r[ids, 0].copy()
            r[ids, 1] = np.minimum(r[ids, 1] + 1, sub)
            r[ids, 0] = (ptr_local + 1) % sub
            self.last_index[ids] = ptr_local + self._offset[ids]
            self._lengths[ids] = r[ids, 1]
            self._child_last[ids, 0] = ptr_local
        self._ep_on_device = True
        self._version += 1
        self._dev_add_version = self._version
        return out_i[0], ep_rew_out, out_i[1], out_i[2]

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
            # "keep" holds the tensor alive so that its address cannot be handed to another
            # tensor while this record is valid (is_canonical compares addresses)
            self._canon = {"ptr": out.data_ptr(), "n": out.numel(), "E": E, "T": T,
                           "version": self._version, "keep": out}
            return out
        sizes = torch.from_numpy(self._rings[:, 1].copy()).to(dev)
        start = torch.from_numpy(self._rings[:, 0].copy()).to(dev)
        off = torch.from_numpy(self._offset).to(dev)
        total = int(self._rings[:, 1].sum())
        k = torch.repeat_interleave(torch.arange(self.buffer_num, device=dev), sizes,
                                    output_size=total)
        first = torch.cumsum(sizes, 0) - sizes
        j = torch.arange(total, device=dev) - first[k]
        out = (start[k] + j) % torch.clamp(sizes[k], min=1) + off[k]
        # remembered so that the return estimators can treat a batch that IS this listing as
        # runs of consecutive slots instead of N arbitrary indices (listing_runs)
        self._listing = {"ptr": out.data_ptr(), "n": out.numel(), "version": self._version,
                         "start": start.contiguous(), "first": first.contiguous(), "keep": out}
        return out

    def listing_runs(self, indices: Any) -> Optional[Dict[str, torch.Tensor]]:
        """If ``indices`` is the tensor the last ``sample_indices(0)`` returned for the current
        contents (ragged or wrapped sub-buffers), the run table of that listing: sub-buffer k
        sits at batch positions ``first[k] + j`` with buffer index
        ``(start[k] + j) % lengths[k] + offset[k]``."""
        c = getattr(self, "_listing", None)
        if (c is not None and isinstance(indices, torch.Tensor)
                and c["version"] == self._version and indices.data_ptr() == c["ptr"]
                and indices.numel() == c["n"]):
            return c
        return None

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
