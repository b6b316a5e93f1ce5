"""``to_numpy`` / ``to_torch`` / ``to_torch_as`` with the semantics of
tianshou/data/utils/converter.py:13-78 (the reference applies them liberally at the boundary of
the hot path: a2c.py:91-93, :115-116; prio.py:86).  Here they are conveniences for callers:
the device-resident path itself does not round-trip through the host."""
from __future__ import annotations

from numbers import Number
from typing import Any, Optional, Union

import numpy as np
import torch

from .batch import Batch, _to_leaf


def to_numpy(x: Any) -> Union[Batch, np.ndarray]:
    if isinstance(x, torch.Tensor):
        return x.detach().cpu().numpy()
    if isinstance(x, np.ndarray):
        return x
    if isinstance(x, (np.number, np.bool_, Number)):
        return np.asanyarray(x)
    if x is None:
        return np.array(None, dtype=object)
    if isinstance(x, (dict, Batch)):
        out = Batch(x).copy() if isinstance(x, Batch) else Batch(x)
        return out.to_numpy()
    if isinstance(x, (list, tuple)):
        return to_numpy(_to_leaf(x))
    return np.asanyarray(x)


def to_torch(x: Any, dtype: Optional[torch.dtype] = None,
             device: Union[str, int, torch.device] = "cpu") -> Union[Batch, torch.Tensor]:
    if isinstance(x, np.ndarray) and issubclass(x.dtype.type, (np.bool_, np.number)):
        t = torch.from_numpy(x).to(device)
        return t.type(dtype) if dtype is not None else t
    if isinstance(x, torch.Tensor):
        if dtype is not None:
            x = x.type(dtype)
        return x.to(device)
    if isinstance(x, (np.number, np.bool_, Number)):
        return to_torch(np.asanyarray(x), dtype, device)
    if isinstance(x, (dict, Batch)):
        out = Batch(x).copy() if isinstance(x, Batch) else Batch(x)
        return out.to_torch(dtype, device)
    if isinstance(x, (list, tuple)):
        return to_torch(_to_leaf(x), dtype, device)
    raise TypeError(f"object {x} cannot be converted to torch.")


def to_torch_as(x: Any, y: torch.Tensor) -> Union[Batch, torch.Tensor]:
    """Same as ``to_torch(x, dtype=y.dtype, device=y.device)``."""
    assert isinstance(y, torch.Tensor)
    return to_torch(x, dtype=y.dtype, device=y.device)
