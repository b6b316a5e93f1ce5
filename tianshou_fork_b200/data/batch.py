"""Minimal ``Batch`` container: the subset of ``tianshou.data.Batch`` the hot path uses.

The reference class (tianshou/data/batch.py:1-912) is a generic nested dict-of-arrays with
object/heterogeneous support; SURVEY.md 2.1 #8 marks that machinery out of scope.  What the
return / replay path needs is: attribute and key access, fancy indexing of every leaf
(``batch[indices]``), scatter assignment (``storage[ptrs] = batch``), ``len``, ``keys``,
``pop``, ``is_empty``, nested batches (``obs`` dicts, ``info``), storage allocation
(``_create_value``, batch.py:94-131), ``split`` (batch.py:896-912), and the small conveniences
callers lean on (``update``, ``copy``, ``shape``, ``cat`` / ``cat_`` of same-keyed batches,
``+ * /``).  Leaves may be NumPy
arrays (host) or torch tensors (device); device-resident buffers keep torch CUDA leaves.
"""
from __future__ import annotations

from numbers import Number
from typing import Any, Dict, Iterator, List, Optional, Union

import numpy as np
import torch

Leaf = Union[np.ndarray, torch.Tensor, "Batch"]


def _to_leaf(v: Any) -> Any:
    if isinstance(v, (np.ndarray, torch.Tensor, Batch)):
        return v
    if isinstance(v, dict):
        return Batch(v)
    if isinstance(v, (Number, np.number, np.bool_)):
        return np.asanyarray(v)
    if isinstance(v, (list, tuple)):
        if len(v) > 0 and all(isinstance(e, (dict, Batch)) for e in v):
            return Batch.stack([Batch(e) if isinstance(e, dict) else e for e in v])
        return np.asanyarray(v)
    return v


def _index_for(leaf: Any, index: Any) -> Any:
    """Adapt an index (numpy / torch / int / slice) to the leaf's array library."""
    if isinstance(leaf, torch.Tensor):
        if isinstance(index, np.ndarray):
            return torch.from_numpy(index).to(leaf.device)
        if isinstance(index, torch.Tensor) and index.device != leaf.device:
            return index.to(leaf.device)
        if isinstance(index, list):
            return torch.as_tensor(index, device=leaf.device)
        return index
    if isinstance(index, torch.Tensor):
        return index.cpu().numpy()
    return index


class Batch:
    """Dict of arrays with attribute access; every leaf shares the first dimension."""

    def __init__(self, batch_dict: Optional[Union[dict, "Batch"]] = None, **kwargs: Any) -> None:
        if batch_dict is not None:
            items = batch_dict.items()
            for k, v in items:
                self.__dict__[k] = _to_leaf(v)
        for k, v in kwargs.items():
            self.__dict__[k] = _to_leaf(v)

    # ---- mapping protocol ------------------------------------------------------------------
    def __setattr__(self, key: str, value: Any) -> None:
        self.__dict__[key] = _to_leaf(value)

    def __getattr__(self, key: str) -> Any:  # only reached when the key is absent
        raise AttributeError(key)

    def __contains__(self, key: str) -> bool:
        return key in self.__dict__

    def keys(self):
        return self.__dict__.keys()

    def values(self):
        return self.__dict__.values()

    def items(self):
        return self.__dict__.items()

    def get(self, key: str, default: Any = None) -> Any:
        return self.__dict__.get(key, default)

    def pop(self, key: str, *default: Any) -> Any:
        return self.__dict__.pop(key, *default)

    def is_empty(self, recurse: bool = False) -> bool:
        if len(self.__dict__) == 0:
            return True
        if not recurse:
            return False
        return all(isinstance(v, Batch) and v.is_empty(True) for v in self.values())

    def __getstate__(self) -> Dict[str, Any]:
        return dict(self.__dict__)

    def __setstate__(self, state: Dict[str, Any]) -> None:
        self.__dict__.update(state)

    def __repr__(self) -> str:
        inner = ", ".join(f"{k}: {getattr(v, 'shape', v)!r}" for k, v in self.items())
        return f"Batch({inner})"

    # ---- indexing --------------------------------------------------------------------------
    def __getitem__(self, index: Any) -> Any:
        if isinstance(index, str):
            return self.__dict__[index]
        if len(self.__dict__) == 0:
            raise IndexError("Cannot access item from empty Batch object.")
        out = Batch()
        for k, v in self.items():
            if isinstance(v, Batch) and v.is_empty():
                out.__dict__[k] = Batch()
            else:
                out.__dict__[k] = v[_index_for(v, index)]
        return out

    def __setitem__(self, index: Any, value: Any) -> None:
        if isinstance(index, str):
            self.__dict__[index] = _to_leaf(value)
            return
        if not isinstance(value, Batch):
            raise ValueError("Batch does not support tensor assignment; use a Batch value")
        if not set(value.keys()).issubset(self.keys()):
            raise ValueError("Creating keys is not supported by item assignment.")
        for k, v in self.items():
            if k not in value:  # reference zero-fills missing keys (batch.py:300-310)
                if isinstance(v, Batch):
                    continue
                v[_index_for(v, index)] = 0
                continue
            src = value[k]
            if isinstance(v, Batch):
                if not (isinstance(src, Batch) and src.is_empty()):
                    v[index] = src
                continue
            idx = _index_for(v, index)
            if isinstance(v, torch.Tensor):
                if not isinstance(src, torch.Tensor):
                    src = torch.as_tensor(np.asarray(src))
                v[idx] = src.to(device=v.device, dtype=v.dtype)
            else:
                if isinstance(src, torch.Tensor):
                    src = src.detach().cpu().numpy()
                v[idx] = src

    def __len__(self) -> int:
        for v in self.values():
            if isinstance(v, Batch):
                if v.is_empty(recurse=True):
                    continue
                return len(v)
            if getattr(v, "ndim", 0) > 0:
                return len(v)
            raise TypeError(f"Object {v} in {self} has no len()")
        return 0

    def __iter__(self) -> Iterator["Batch"]:
        for i in range(len(self)):
            yield self[i]

    # ---- transforms ------------------------------------------------------------------------
    def to_torch(self, dtype: Optional[torch.dtype] = None,
                 device: Union[str, torch.device] = "cpu") -> "Batch":
        for k, v in self.items():
            if isinstance(v, Batch):
                v.to_torch(dtype, device)
            elif isinstance(v, np.ndarray):
                t = torch.from_numpy(np.ascontiguousarray(v)).to(device)
                self.__dict__[k] = t.to(dtype) if dtype is not None and t.is_floating_point() else t
            elif isinstance(v, torch.Tensor):
                self.__dict__[k] = v.to(device)
        return self

    def to_numpy(self) -> "Batch":
        for k, v in self.items():
            if isinstance(v, Batch):
                v.to_numpy()
            elif isinstance(v, torch.Tensor):
                self.__dict__[k] = v.detach().cpu().numpy()
        return self

    @staticmethod
    def stack(batches: List["Batch"], axis: int = 0) -> "Batch":
        out = Batch()
        if not batches:
            return out
        for k in batches[0].keys():
            vs = [b[k] for b in batches]
            if isinstance(vs[0], Batch):
                out.__dict__[k] = Batch() if vs[0].is_empty() else Batch.stack(vs, axis)
            elif isinstance(vs[0], torch.Tensor):
                out.__dict__[k] = torch.stack(vs, dim=axis)
            else:
                out.__dict__[k] = np.stack([np.asanyarray(v) for v in vs], axis=axis)
        return out

    def update(self, batch: Optional[Union[dict, "Batch"]] = None, **kwargs: Any) -> None:
        """Set / overwrite keys like ``dict.update`` (batch.py:816-825)."""
        if batch is not None:
            for k, v in (batch.items() if isinstance(batch, (dict, Batch)) else batch):
                self.__dict__[k] = _to_leaf(v)
        for k, v in kwargs.items():
            self.__dict__[k] = _to_leaf(v)

    def copy(self) -> "Batch":
        out = Batch()
        for k, v in self.items():
            if isinstance(v, Batch):
                out.__dict__[k] = v.copy()
            elif isinstance(v, torch.Tensor):
                out.__dict__[k] = v.clone()
            else:
                out.__dict__[k] = np.array(v, copy=True)
        return out

    @property
    def shape(self) -> List[int]:
        """Common leading shape of the leaves (batch.py:878-894)."""
        if self.is_empty():
            return []
        shapes = []
        for v in self.values():
            if isinstance(v, Batch):
                if not v.is_empty(recurse=True):
                    shapes.append(list(v.shape))
            else:
                shapes.append(list(getattr(v, "shape", [])))
        if not shapes:
            return []
        n = min(len(sh) for sh in shapes)
        out = []
        for d in range(n):
            if len({sh[d] for sh in shapes}) != 1:
                break
            out.append(shapes[0][d])
        return out

    @staticmethod
    def cat(batches: List[Union[dict, "Batch"]]) -> "Batch":
        """Concatenate along the first axis (batch.py:664-703); the batches share their keys."""
        bs = [b if isinstance(b, Batch) else Batch(b) for b in batches]
        bs = [b for b in bs if not b.is_empty(recurse=True)]
        out = Batch()
        if not bs:
            return out
        for k in bs[0].keys():
            vs = [b[k] for b in bs]
            if isinstance(vs[0], Batch):
                out.__dict__[k] = Batch() if vs[0].is_empty() else Batch.cat(vs)
            elif isinstance(vs[0], torch.Tensor):
                out.__dict__[k] = torch.cat(vs, dim=0)
            else:
                out.__dict__[k] = np.concatenate([np.asanyarray(v) for v in vs], axis=0)
        return out

    def cat_(self, batches: Union["Batch", List[Union[dict, "Batch"]]]) -> None:
        if isinstance(batches, (Batch, dict)):
            batches = [batches]
        merged = Batch.cat(([self] if not self.is_empty() else []) + list(batches))
        self.__dict__.clear()
        self.__dict__.update(merged.__dict__)

    def _apply(self, fn) -> "Batch":
        for k, v in self.items():
            if isinstance(v, Batch):
                if not v.is_empty():
                    v._apply(fn)
            else:
                self.__dict__[k] = fn(k, v)
        return self

    def __iadd__(self, other: Union["Batch", Number]) -> "Batch":
        if isinstance(other, Batch):
            return self._apply(lambda k, v: v + other[k])
        return self._apply(lambda k, v: v + other)

    def __add__(self, other: Union["Batch", Number]) -> "Batch":
        return self.copy().__iadd__(other)

    def __imul__(self, value: Number) -> "Batch":
        return self._apply(lambda k, v: v * value)

    def __mul__(self, value: Number) -> "Batch":
        return self.copy().__imul__(value)

    def __itruediv__(self, value: Number) -> "Batch":
        return self._apply(lambda k, v: v / value)

    def __truediv__(self, value: Number) -> "Batch":
        return self.copy().__itruediv__(value)

    def split(self, size: int, shuffle: bool = True, merge_last: bool = False
              ) -> Iterator["Batch"]:
        """Minibatch iterator (tianshou/data/batch.py:896-912): host ``np.random.permutation``
        when shuffling, the short tail folded into the previous chunk with ``merge_last``."""
        length = len(self)
        if size == -1:
            size = length
        assert size >= 1
        indices = np.random.permutation(length) if shuffle else np.arange(length)
        merge_last = merge_last and length % size > 0
        for idx in range(0, length, size):
            if merge_last and idx + size + size >= length:
                yield self[indices[idx:]]
                break
            yield self[indices[idx:idx + size]]


def create_value(inst: Any, size: int, stack: bool = True, device: Optional[torch.device] = None
                 ) -> Any:
    """Allocate storage of ``size`` slots shaped like ``inst`` (batch.py:94-131).

    ``stack=False`` means ``inst`` already carries a leading batch dimension that is replaced
    by ``size``.  With ``device`` set, leaves are torch tensors on that device; ``rew`` is
    widened to float64 and flags to bool by the caller exactly as the reference does
    (manager.py:150-153) before this is called.
    """
    if isinstance(inst, Batch):
        out = Batch()
        for k, v in inst.items():
            out.__dict__[k] = create_value(v, size, stack, device)
        return out
    if isinstance(inst, torch.Tensor):
        shape = (size, *inst.shape) if stack else (size, *inst.shape[1:])
        return torch.zeros(shape, dtype=inst.dtype, device=device or inst.device)
    arr = np.asanyarray(inst)
    shape = (size, *arr.shape) if stack else (size, *arr.shape[1:])
    if device is not None:
        if arr.dtype == object:
            raise TypeError("object arrays cannot live in a device-resident buffer")
        return torch.zeros(shape, dtype=torch.from_numpy(np.zeros(0, arr.dtype)).dtype,
                           device=device)
    return np.zeros(shape, dtype=arr.dtype)
