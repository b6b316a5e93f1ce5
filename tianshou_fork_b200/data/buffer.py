"""Device-resident replay buffers -- drop-ins for ``tianshou.data.ReplayBuffer``,
``ReplayBufferManager`` and ``VectorReplayBuffer``
(tianshou/data/buffer/base.py:11-389, manager.py:10-192, vecbuf.py:15-37).

Storage (``obs/act/rew/terminated/truncated/done/obs_next/...``) lives in HBM as torch CUDA
tensors inside a :class:`Batch`; ``prev`` / ``next`` / ``unfinished_index`` run as integer
kernels (``csrc/index.cu``).  Ring bookkeeping (``_index``, ``_size``, ``last_index``,
``_lengths``, episode statistics) stays on the host -- it is a handful of integers per
sub-buffer that ``add`` needs synchronously for its return value -- and is mirrored to the
device lazily, before the first kernel that needs it.

Index arguments may be Python ints, lists, NumPy arrays or CUDA tensors; NumPy/int in ->
NumPy/int out, CUDA tensor in -> CUDA tensor out.  ``sample_indices`` / ``sample`` return CUDA
tensors so that the whole update stays on the device.

Out of scope here: HDF5 save/load (no h5py in this image).
"""
from __future__ import annotations

import ctypes
from typing import Any, Dict, List, Optional, Tuple, Union

import numpy as np
import torch

from .. import _lib as L
from .batch import Batch, create_value

IndexLike = Union[int, List[int], np.ndarray, torch.Tensor]


def _is_scalar(index: Any) -> bool:
    return isinstance(index, (int, np.integer)) or (
        isinstance(index, np.ndarray) and index.ndim == 0)


_NP_OF_TORCH = {torch.float32: np.float32, torch.float64: np.float64, torch.int64: np.int64,
                torch.int32: np.int32, torch.uint8: np.uint8, torch.bool: np.bool_,
                torch.int16: np.int16, torch.int8: np.int8, torch.float16: np.float16}


class ReplayBuffer:
    """Circular buffer with device-resident storage (tianshou/data/buffer/base.py:11-389)."""

    _reserved_keys = (
        "obs", "act", "rew", "terminated", "truncated", "done", "obs_next", "info", "policy"
    )
    _input_keys = (
        "obs", "act", "rew", "terminated", "truncated", "obs_next", "info", "policy"
    )

    def __init__(
        self,
        size: int,
        stack_num: int = 1,
        ignore_obs_next: bool = False,
        save_only_last_obs: bool = False,
        sample_avail: bool = False,
        device: Union[str, torch.device, None] = None,
        rew_dtype: torch.dtype = torch.float64,
        **kwargs: Any,
    ) -> None:
        L.lib()  # no extension or no GPU -> fail here, loudly
        self.options: Dict[str, Any] = {
            "stack_num": stack_num,
            "ignore_obs_next": ignore_obs_next,
            "save_only_last_obs": save_only_last_obs,
            "sample_avail": sample_avail,
        }
        self.device = torch.device(device if device is not None else
                                   f"cuda:{torch.cuda.current_device()}")
        self.rew_dtype = rew_dtype
        self.maxsize = int(size)
        assert stack_num > 0, "stack_num should be greater than 0"
        self.stack_num = stack_num
        self._save_obs_next = not ignore_obs_next
        self._save_only_last_obs = save_only_last_obs
        self._sample_avail = sample_avail
        self._meta: Batch = Batch()
        # ring state [index, size]; a manager re-points this at a row of its own table
        self._ring = np.zeros(2, dtype=np.int64)
        self._ep = {"rew": 0.0, "len": 0, "idx": 0}
        self._version = 0
        self._parent: Optional["ReplayBuffer"] = None  # set by a manager on its children
        self._dev_state: Optional[Dict[str, torch.Tensor]] = None
        self._dev_state_version = -1
        self.reset()

    # ---- ring state ------------------------------------------------------------------------
    @property
    def _index(self) -> int:
        return int(self._ring[0])

    @_index.setter
    def _index(self, v: int) -> None:
        self._ring[0] = v

    @property
    def _size(self) -> int:
        return int(self._ring[1])

    @_size.setter
    def _size(self, v: int) -> None:
        self._ring[1] = v

    def __len__(self) -> int:
        return self._size

    def __repr__(self) -> str:
        return self.__class__.__name__ + self._meta.__repr__()[5:]

    def __getattr__(self, key: str) -> Any:
        """``buf.rew`` etc. hand out the live device storage (base.py:74-79)."""
        try:
            return self.__dict__["_meta"][key]
        except KeyError as exception:
            raise AttributeError(key) from exception

    def __setattr__(self, key: str, value: Any) -> None:
        assert key not in self._reserved_keys, \
            "key '{}' is reserved and cannot be assigned".format(key)
        super().__setattr__(key, value)

    def __getstate__(self) -> Dict[str, Any]:
        """Pickle as host arrays (base.py:81-87; test/base/test_buffer.py:597-632)."""
        if getattr(self, "_ep_on_device", False):
            self._pull_episode_stats()  # manager: running episode statistics back to the host
        state = dict(self.__dict__)
        if state.get("_parent") is not None:
            state["_meta"] = Batch()  # a sub-buffer's storage is a view: the manager restores it
        else:
            state["_meta"] = _batch_to_numpy_copy(Batch(self._meta))
        # caches that hold device addresses or ctypes arrays are rebuilt on demand
        for key in ("_dev_state", "_dev_add", "_dev_add_keys", "_canon", "_listing",
                    "_cached_weight"):
            if key in state:
                state[key] = None
        for key in ("_pin", "_host_io", "_host_io_upd", "_upd_ev"):  # staging of host-array calls
            state.pop(key, None)
        state["_dev_state_version"] = -1
        if "_dev_add_version" in state:
            state["_dev_add_version"] = -1
        state["device"] = str(self.device)
        return state

    def __setstate__(self, state: Dict[str, Any]) -> None:
        self.__dict__.update(state)
        self.__dict__["device"] = torch.device(state["device"])
        self.__dict__["_meta"] = state["_meta"].to_torch(device=self.device)

    @classmethod
    def from_data(cls, obs: Any, act: Any, rew: Any, terminated: Any, truncated: Any,
                  done: Any, obs_next: Any) -> "ReplayBuffer":
        """A full buffer over existing arrays (offline datasets; base.py:108-132).  The arrays
        are uploaded once; they need not be HDF5 datasets."""
        size = len(obs)
        assert all(len(d) == size for d in [obs, act, rew, terminated, truncated, done,
                                            obs_next]), \
            "Lengths of all hdf5 datasets need to be equal."
        buf = cls(size)
        if size == 0:
            return buf
        as_arr = lambda x: x if isinstance(x, torch.Tensor) else np.asarray(x)  # noqa: E731
        batch = Batch(obs=as_arr(obs), act=as_arr(act), rew=as_arr(rew),
                      terminated=as_arr(terminated), truncated=as_arr(truncated),
                      done=as_arr(done), obs_next=as_arr(obs_next))
        buf.set_batch(batch)
        for k in ("terminated", "truncated", "done"):
            buf._meta.__dict__[k] = buf._meta[k].to(torch.bool)
        buf._size = size
        return buf

    def reset(self, keep_statistics: bool = False) -> None:
        """Clear all the data in replay buffer and episode statistics (base.py:134-139)."""
        if self.__dict__.get("_parent") is not None:
            self.last_index[:] = 0  # a sub-buffer's row of its manager's table: keep the alias
        else:
            self.last_index = np.array([0])
        self._ring[:] = 0
        if not keep_statistics:
            self._ep = {"rew": 0.0, "len": 0, "idx": 0}
        self._version += 1

    def set_batch(self, batch: Batch) -> None:
        """Manually choose the batch you want the ReplayBuffer to manage (base.py:141-146)."""
        assert len(batch) == self.maxsize and set(batch.keys()).issubset(
            self._reserved_keys
        ), "Input batch doesn't meet ReplayBuffer's data form requirement."
        self._meta = batch.to_torch(device=self.device)
        self._version += 1

    # ---- device mirrors of the bookkeeping ---------------------------------------------------
    def _layout(self) -> Tuple[np.ndarray, np.ndarray, np.ndarray, int]:
        """(extend_offset[B+1], last_index[B], lengths[B], sub_size) for the index kernels.

        A single ring takes ``%`` by its CURRENT size (base.py:158-168), which is the manager
        formula (manager.py:267-276) with one sub-buffer of capacity ``size``.  The current
        size travels in the device arrays only (``offset[1]``, ``lengths[0]``); the by-value
        ``sub_size`` is the constant capacity, so a captured graph stays valid while the ring
        fills (the kernels never divide by it when B == 1)."""
        size = max(self._size, 1)
        return (np.array([0, size], dtype=np.int64),
                np.asarray(self.last_index, dtype=np.int64).reshape(1),
                np.array([self._size], dtype=np.int64), max(self.maxsize, 1))

    def _state(self) -> Dict[str, torch.Tensor]:
        version = self._version + (self._parent._version if self._parent is not None else 0)
        if self._dev_state is None or self._dev_state_version != version:
            off, last, lens, sub = self._layout()
            packed = np.concatenate([off, last, lens]).astype(np.int64)
            old = self._dev_state["packed"] if self._dev_state is not None else None
            if old is not None and old.numel() == packed.size:
                # refresh in place (stream-ordered): kernels already captured into a CUDA graph
                # keep reading the bookkeeping at the same addresses
                t = old
                t.copy_(torch.from_numpy(packed))
            else:
                t = torch.from_numpy(packed).to(self.device)
            b = last.size
            self._dev_state = {"offset": t[: b + 1], "last_index": t[b + 1: 2 * b + 1],
                               "lengths": t[2 * b + 1:], "B": b, "sub_size": sub,
                               "packed": t}
            # raw addresses for the hot calls (the packed tensor stays alive in the dict)
            p0 = t.data_ptr()
            self._dev_state["ptrs"] = (p0, p0 + 8 * (b + 1), p0 + 8 * (2 * b + 1))
            self._dev_state_version = version
        return self._dev_state

    def _done_u8(self) -> torch.Tensor:
        return L.as_u8(self._meta["done"])

    def _neighbour(self, index: IndexLike, fn_name: str):
        st = self._state()
        if (isinstance(index, torch.Tensor) and index.is_cuda and index.dtype == torch.int64
                and index.is_contiguous()):
            # device tensor in -> device tensor out: nothing but the launch
            sh = L.shim()
            if sh is not None:
                return sh.neighbour(index, st["packed"], st["B"], st["sub_size"],
                                    self._meta.__dict__["done"], fn_name == "ts_next_index")
            out = torch.empty_like(index)
            n = index.numel()
            if n:
                off, last, lens = st["ptrs"]
                L.check(getattr(L.lib(), fn_name)(
                    index.data_ptr(), n, off, st["B"], st["sub_size"],
                    self._meta.__dict__["done"].data_ptr(), last, lens, out.data_ptr(),
                    L.stream()), fn_name)
            return out
        scalar = _is_scalar(index)
        if isinstance(index, torch.Tensor):
            idx = index.to(device=self.device, dtype=torch.int64)
            shape = idx.shape
        else:
            arr = np.asarray(index, dtype=np.int64)
            shape = arr.shape
            idx = torch.from_numpy(np.ascontiguousarray(arr)).to(self.device)
        idx = idx.reshape(-1).contiguous()
        out = torch.empty_like(idx)
        if idx.numel():
            fn = getattr(L.lib(), fn_name)
            L.check(fn(L.ptr(idx), idx.numel(), L.ptr(st["offset"]), st["B"], st["sub_size"],
                       L.ptr(self._done_u8()), L.ptr(st["last_index"]), L.ptr(st["lengths"]),
                       L.ptr(out), L.stream()), fn_name)
        out = out.reshape(shape)
        if isinstance(index, torch.Tensor):
            return out
        res = out.cpu().numpy()
        return res[()] if scalar else res

    def prev(self, index: IndexLike):
        """Index of the previous transition; unchanged at an episode start (base.py:153-160)."""
        return self._neighbour(index, "ts_prev_index")

    def next(self, index: IndexLike):
        """Index of the next transition; unchanged at an episode end (base.py:162-168)."""
        return self._neighbour(index, "ts_next_index")

    def unfinished_index(self) -> np.ndarray:
        """Index of the unfinished episode's last step (base.py:148-151).  Host array: it has
        at most one entry per sub-buffer and the bookkeeping is host-side; the ``done`` bit
        costs one small D2H read."""
        u = self.unfinished_index_device()
        u = u[u >= 0]
        return u.cpu().numpy()

    def unfinished_index_device(self) -> torch.Tensor:
        """int64[B]: unfinished index per sub-buffer or -1 (device, no sync)."""
        st = self._state()
        out = torch.empty(st["B"], dtype=torch.int64, device=self.device)
        if "done" not in self._meta:
            return out.fill_(-1)
        L.check(L.lib().ts_unfinished_index(
            L.ptr(self._done_u8()), L.ptr(st["last_index"]), L.ptr(st["lengths"]), st["B"],
            L.ptr(out), L.stream()), "ts_unfinished_index")
        return out

    # ---- writes ----------------------------------------------------------------------------
    def update(self, buffer: "ReplayBuffer") -> np.ndarray:
        """Move the data from the given buffer to current buffer (base.py:170-193)."""
        if len(buffer) == 0 or self.maxsize == 0:
            return np.array([], int)
        stack_num, buffer.stack_num = buffer.stack_num, 1
        from_indices = buffer.sample_indices(0)
        buffer.stack_num = stack_num
        n = len(from_indices)
        if n == 0:
            return np.array([], int)
        to_indices = (self._index + np.arange(n)) % self.maxsize
        self.last_index[0] = to_indices[-1]
        self._index = int((self._index + n) % self.maxsize)
        self._size = min(self._size + n, self.maxsize)
        self._version += 1
        if self._meta.is_empty():
            self._meta = create_value(buffer._meta, self.maxsize, stack=False,
                                      device=self.device)
        self._meta[to_indices] = buffer._meta[from_indices]
        return to_indices

    def _add_index(self, rew: Union[float, np.ndarray], done: bool
                   ) -> Tuple[int, Union[float, np.ndarray], int, int]:
        """Advance the ring after adding one transition (base.py:195-214)."""
        self.last_index[0] = ptr = self._index
        self._size = min(self._size + 1, self.maxsize)
        self._index = (self._index + 1) % self.maxsize
        self._version += 1
        ep = self._ep
        ep["rew"] = ep["rew"] + rew
        ep["len"] += 1
        if done:
            result = ptr, ep["rew"], ep["len"], ep["idx"]
            ep["rew"], ep["len"], ep["idx"] = 0.0, 0, self._index
            return result
        return ptr, ep["rew"] * 0.0, 0, ep["idx"]

    def _prepare_add(self, batch: Batch, keys, stacked: bool) -> Batch:
        new_batch = Batch()
        for key in set(keys).intersection(batch.keys()):
            new_batch.__dict__[key] = batch[key]
        batch = new_batch
        term, trunc = _host(batch.terminated), _host(batch.truncated)
        batch.__dict__["done"] = np.logical_or(term, trunc)
        assert set(["obs", "act", "rew", "terminated", "truncated", "done"]).issubset(
            batch.keys())
        if self._save_only_last_obs:
            batch.obs = batch.obs[:, -1] if stacked else batch.obs[-1]
        if not self._save_obs_next:
            batch.pop("obs_next", None)
        elif self._save_only_last_obs:
            batch.obs_next = batch.obs_next[:, -1] if stacked else batch.obs_next[-1]
        return batch

    def _widen(self, batch: Batch) -> Batch:
        """rew -> float, flags -> bool before the first allocation (base.py:264-267)."""
        batch.rew = _astype(batch.rew, self.rew_dtype)
        for k in ("done", "terminated", "truncated"):
            batch.__dict__[k] = _astype(batch[k], torch.bool)
        return batch

    def add(self, batch: Batch, buffer_ids: Optional[Union[np.ndarray, List[int]]] = None
            ) -> Tuple[np.ndarray, np.ndarray, np.ndarray, np.ndarray]:
        """Add a batch of data into replay buffer (base.py:216-274)."""
        stacked_batch = buffer_ids is not None
        batch = self._prepare_add(batch, self._input_keys, stacked_batch)
        if stacked_batch:
            assert len(batch) == 1
            rew, done = _host(batch.rew)[0], bool(batch.done[0])
        else:
            rew, done = _host(batch.rew), bool(batch.done)
        ptr, ep_rew, ep_len, ep_idx = list(
            map(lambda x: np.array([x]), self._add_index(rew, done)))
        batch = self._widen(batch)
        if self._meta.is_empty():
            self._meta = create_value(batch, self.maxsize, stack=not stacked_batch,
                                      device=self.device)
        elif not set(batch.keys()).issubset(self._meta.keys()):
            _alloc_missing(self._meta, batch, self.maxsize, not stacked_batch, self.device)
        rows = batch if stacked_batch else _unsqueeze0(batch)
        if not self._store_packed(ptr, rows):
            self._meta[ptr] = rows
        return ptr, ep_rew, ep_len, ep_idx

    def _store_packed(self, ptrs: np.ndarray, batch: Batch) -> bool:
        """``self._meta[ptrs] = batch`` (manager.py:160) for a step that arrives as NumPy
        arrays: the slot indices and the rows of every key travel to the device as ONE packed
        buffer (one copy instead of two per key) and are written by one ``ts_scatter_fields``
        launch.  Returns False (nothing written) for anything it does not cover -- nested
        non-empty batches, device tensors, a shape that does not match the storage."""
        m = int(ptrs.size)
        if m == 0:
            return True
        if m > 1 and np.unique(ptrs).size != m:
            return False  # a slot written twice in one call: keep NumPy's last-writer order
        meta = self._meta.__dict__
        fields = []
        for k, v in batch.items():
            dst = meta.get(k)
            if isinstance(v, Batch):
                if v.is_empty() and isinstance(dst, Batch):
                    continue
                return False
            if not (isinstance(v, np.ndarray) and isinstance(dst, torch.Tensor) and dst.is_cuda
                    and dst.is_contiguous() and v.shape == (m,) + tuple(dst.shape[1:])):
                return False
            np_dtype = _NP_OF_TORCH.get(dst.dtype)
            if np_dtype is None:
                return False
            fields.append((dst, np.ascontiguousarray(v, dtype=np_dtype)))
        if len(fields) > 8 or set(batch.keys()) != set(k for k in meta if k in batch):
            return False
        for k in meta:  # the reference zero-fills stored keys the step does not carry
            if k not in batch and not isinstance(meta[k], Batch):
                return False
        offsets, total = [], 8 * m
        for dst, arr in fields:
            offsets.append(total)
            total += (arr.nbytes + 7) & ~7
        packed = np.empty(total, dtype=np.uint8)
        packed[:8 * m] = np.ascontiguousarray(ptrs, dtype=np.int64).view(np.uint8)
        for off, (dst, arr) in zip(offsets, fields):
            packed[off:off + arr.nbytes] = arr.reshape(-1).view(np.uint8)
        dev_packed = torch.from_numpy(packed).to(self.device)
        base = dev_packed.data_ptr()
        nf = len(fields)
        src, dsts, rb = ((ctypes.c_void_p * nf)(), (ctypes.c_void_p * nf)(),
                         (ctypes.c_int64 * nf)())
        for j, (off, (dst, arr)) in enumerate(zip(offsets, fields)):
            src[j], dsts[j], rb[j] = base + off, dst.data_ptr(), arr.nbytes // m
        L.check(L.lib().ts_scatter_fields(nf, src, dsts, rb, self.maxsize, base, m, L.stream()),
                "ts_scatter_fields")
        return True

    # ---- sampling / gathers ------------------------------------------------------------------
    def _all_indices_host(self) -> np.ndarray:
        return np.concatenate([np.arange(self._index, self._size), np.arange(self._index)])

    def sample_indices(self, batch_size: int) -> torch.Tensor:
        """Random indices (host ``np.random`` stream, same draws as the reference), all valid
        indices in chronological order for 0, empty for negative (base.py:276-305)."""
        if self.stack_num == 1 or not self._sample_avail:
            if batch_size > 0:
                idx = np.random.choice(self._size, batch_size)
            elif batch_size == 0:
                idx = self._all_indices_host()
            else:
                idx = np.array([], int)
        else:
            if batch_size < 0:
                idx = np.array([], int)
            else:
                all_indices = prev_indices = self._all_indices_host()
                for _ in range(self.stack_num - 2):
                    prev_indices = self.prev(prev_indices)
                all_indices = all_indices[prev_indices != self.prev(prev_indices)]
                idx = np.random.choice(all_indices, batch_size) if batch_size > 0 \
                    else all_indices
        return torch.from_numpy(np.asarray(idx, dtype=np.int64)).to(self.device)

    def sample(self, batch_size: int) -> Tuple[Batch, torch.Tensor]:
        """Random batch and its indices; all data for ``batch_size == 0`` (base.py:307-315)."""
        indices = self.sample_indices(batch_size)
        return self[indices], indices

    def get(self, index: IndexLike, key: str, default_value: Any = None,
            stack_num: Optional[int] = None):
        """Stacked gather ``[obs[t-n+1], ..., obs[t]]`` through ``prev`` (base.py:317-358)."""
        if key not in self._meta and default_value is not None:
            return default_value
        val = self._meta[key]
        if isinstance(val, Batch) and val.is_empty():
            return Batch()  # what the reference's loop ends in (IndexError -> Batch(), :355-358)
        if stack_num is None:
            stack_num = self.stack_num
        try:
            if stack_num == 1:
                return _take(val, index)
            if isinstance(val, torch.Tensor) and val.is_cuda and stack_num <= 16:
                return self._stack_gather(val, index, stack_num)[0]
            stack: List[Any] = []
            indices = self._as_device_index(index)
            for _ in range(stack_num):
                stack = [_take(val, indices)] + stack
                indices = self.prev(indices)
            if isinstance(val, Batch):
                return Batch.stack(stack, axis=indices.ndim)
            return torch.stack(stack, dim=indices.ndim)
        except IndexError as exception:
            if not (isinstance(val, Batch) and val.is_empty()):
                raise exception
            return Batch()

    def _stack_gather(self, val: torch.Tensor, index: IndexLike, stack_num: int,
                      with_next: bool = False):
        """``stack([val[prev^(k-1)(i)], ..., val[i]], axis=i.ndim)`` in one launch
        (``ts_gather_stack``); with ``with_next`` also the stack that ends at ``next(i)``
        (the ``obs_next`` of ``ignore_obs_next`` buffers, base.py:374-378)."""
        if not isinstance(index, torch.Tensor):
            arr = np.asarray(index, dtype=np.int64)
            if arr.size and (arr.max() >= val.shape[0] or arr.min() < -val.shape[0]):
                raise IndexError(f"index {arr.max()} is out of bounds for size {val.shape[0]}")
        indices = self._as_device_index(index)
        shape = tuple(indices.shape)
        idx = indices.reshape(-1).contiguous()
        st = self._state()
        val = val if val.is_contiguous() else val.contiguous()
        row_shape = tuple(val.shape[1:])
        row_bytes = val.stride(0) * val.element_size() if val.ndim > 1 else val.element_size()
        out = torch.empty((idx.numel(), stack_num) + row_shape, dtype=val.dtype,
                          device=val.device)
        out_next = torch.empty_like(out) if with_next else None
        if idx.numel():
            L.check(L.lib().ts_gather_stack(
                L.ptr(val), row_bytes, L.ptr(idx), idx.numel(), stack_num, L.ptr(st["offset"]),
                st["B"], st["sub_size"], L.ptr(self._done_u8()), L.ptr(st["last_index"]),
                L.ptr(st["lengths"]), L.ptr(out), L.ptr(out_next) if with_next else None,
                L.stream()), "ts_gather_stack")
        out = out.reshape(shape + (stack_num,) + row_shape)
        if with_next:
            out_next = out_next.reshape(shape + (stack_num,) + row_shape)
        return out, out_next

    def _take_fields(self, keys: Tuple[str, ...], indices: torch.Tensor) -> List[Any]:
        """``[self._meta[k][indices] for k in keys]`` with ONE launch (``ts_gather_fields``)
        for the keys stored as small-row CUDA tensors; anything else (nested batches, wide
        rows) is indexed on its own.  base.py:379-389 does one fancy-index copy per key."""
        out: List[Any] = [None] * len(keys)
        fused: List[int] = []
        meta = self._meta.__dict__
        for i, k in enumerate(keys):
            v = meta[k]
            if (isinstance(v, torch.Tensor) and v.is_cuda and v.ndim >= 1 and v.shape[0] > 0
                    and v.is_contiguous() and v.stride(0) * v.element_size() <= 256):
                fused.append(i)
            else:
                out[i] = _take(v, indices)
        nf = len(fused)
        if nf < 2 or nf > 8 or indices.ndim == 0 or indices.numel() == 0:
            for i in fused:
                out[i] = _take(meta[keys[i]], indices)
            return out
        if not (indices.is_cuda and indices.dtype == torch.int64 and indices.is_contiguous()):
            indices = indices.to(device=self.device, dtype=torch.int64).contiguous()
        shape = tuple(indices.shape)
        arr_p, arr_i = _FIELD_ARRAYS[nf]
        src, dst, rb = arr_p(), arr_p(), arr_i()
        n_rows = 0
        for j, i in enumerate(fused):
            v = meta[keys[i]]
            n_rows = v.shape[0]
            o = torch.empty(shape + tuple(v.shape[1:]), dtype=v.dtype, device=v.device)
            out[i] = o
            src[j], dst[j] = v.data_ptr(), o.data_ptr()
            rb[j] = v.stride(0) * v.element_size() if v.ndim > 1 else v.element_size()
        L.check(L.lib().ts_gather_fields(nf, src, dst, rb, n_rows, indices.data_ptr(),
                                         indices.numel(), L.stream()), "ts_gather_fields")
        return out

    def _as_device_index(self, index: IndexLike) -> torch.Tensor:
        if isinstance(index, torch.Tensor):
            if index.is_cuda and index.dtype == torch.int64 and index.device == self.device:
                return index  # `.to` with nothing to do still costs a dispatcher round trip
            return index.to(device=self.device, dtype=torch.int64)
        return torch.from_numpy(np.asarray(index, dtype=np.int64)).to(self.device)

    def obs_next_of(self, index: IndexLike) -> Any:
        """The ``obs_next`` entry of ``buffer[index]`` alone (base.py:374-378): what a target
        network reads of the n-step terminal transitions (dqn.py:85-86), without gathering the
        other keys."""
        indices = self._as_device_index(index)
        if self._save_obs_next:
            return self.get(indices, "obs_next", Batch())
        return self.get(self.next(indices), "obs", Batch())

    def __getitem__(self, index: Union[slice, IndexLike]) -> Batch:
        """``buffer[index]``: gather copies of every stored key (base.py:360-389)."""
        if isinstance(index, slice):
            indices = self.sample_indices(0) if index == slice(None) \
                else self._as_device_index(np.arange(self.maxsize)[:len(self)][index])
        else:
            indices = self._as_device_index(index)
        obs_val = self._meta["obs"] if "obs" in self._meta else None
        if (not self._save_obs_next and 1 < self.stack_num <= 16
                and isinstance(obs_val, torch.Tensor) and obs_val.is_cuda):
            # obs and obs[next(indices)] stacks from one pass over k + 1 frames
            obs, obs_next = self._stack_gather(obs_val, indices, self.stack_num, with_next=True)
        else:
            obs = self.get(indices, "obs")
            if self._save_obs_next:
                obs_next = self.get(indices, "obs_next", Batch())
            else:
                obs_next = self.get(self.next(indices), "obs", Batch())
        act, rew, terminated, truncated, done = self._take_fields(
            ("act", "rew", "terminated", "truncated", "done"), indices)
        return Batch(
            obs=obs,
            act=act,
            rew=rew,
            terminated=terminated,
            truncated=truncated,
            done=done,
            obs_next=obs_next,
            info=self.get(indices, "info", Batch()),
            policy=self.get(indices, "policy", Batch()),
        )


# ---- helpers ---------------------------------------------------------------------------------
# ctypes array types of the field table, by field count (creating a type per call is slow)
_FIELD_ARRAYS = {n: (ctypes.c_void_p * n, ctypes.c_int64 * n) for n in range(1, 9)}


def _host(x: Any) -> np.ndarray:
    if isinstance(x, torch.Tensor):
        return x.detach().cpu().numpy()
    return np.asarray(x)


def _astype(x: Any, dtype: torch.dtype) -> Any:
    if isinstance(x, torch.Tensor):
        return x.to(dtype)
    np_dtype = {torch.float64: np.float64, torch.float32: np.float32, torch.bool: np.bool_}[dtype]
    return np.asarray(x).astype(np_dtype)


def _unsqueeze0(batch: Batch) -> Batch:
    out = Batch()
    for k, v in batch.items():
        if isinstance(v, Batch):
            out.__dict__[k] = _unsqueeze0(v) if not v.is_empty() else v
        elif isinstance(v, torch.Tensor):
            out.__dict__[k] = v.unsqueeze(0)
        else:
            out.__dict__[k] = np.asarray(v)[None]
    return out


def _take(val: Any, index: Any) -> Any:
    if isinstance(val, Batch):
        return val[index] if not val.is_empty() else Batch()
    if isinstance(index, np.ndarray):
        index = torch.from_numpy(index).to(val.device)
    elif isinstance(index, list):
        index = torch.as_tensor(index, device=val.device)
    elif isinstance(index, (int, np.integer)):
        if index >= val.shape[0] or index < -val.shape[0]:
            raise IndexError(f"index {index} is out of bounds for size {val.shape[0]}")
    return val[index]


def _alloc_missing(meta: Batch, batch: Batch, size: int, stack: bool, device) -> None:
    """Dynamic keys popping up after the first add (``_alloc_by_keys_diff``, base.py:271-272)."""
    for k, v in batch.items():
        if k not in meta:
            meta.__dict__[k] = create_value(v, size, stack, device)
        elif isinstance(v, Batch) and isinstance(meta[k], Batch):
            if meta[k].is_empty():
                meta.__dict__[k] = create_value(v, size, stack, device)
            else:
                _alloc_missing(meta[k], v, size, stack, device)


def _batch_to_numpy_copy(b: Batch) -> Batch:
    out = Batch()
    for k, v in b.items():
        if isinstance(v, Batch):
            out.__dict__[k] = _batch_to_numpy_copy(v)
        elif isinstance(v, torch.Tensor):
            out.__dict__[k] = v.detach().cpu().numpy()
        else:
            out.__dict__[k] = v
    return out
