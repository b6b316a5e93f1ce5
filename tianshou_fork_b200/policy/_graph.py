"""Shared pieces of the CUDA-graphed update steps (``PPOPolicy(cuda_graph=True)``,
``DQNPolicy(cuda_graph=True)``).

Capturing a training step needs warm-up iterations that run for real (allocator, cuDNN /
cuBLAS plans, lazy optimiser state).  ``StateGuard`` remembers everything those iterations
change -- parameters, module buffers, optimiser state, any extra tensors the caller lists --
and puts it back IN PLACE, so that the captured graph sees the same addresses and the training
trajectory is the one the eager path would have taken.
"""
from __future__ import annotations

from contextlib import contextmanager
from typing import Any, Callable, Iterable, Iterator, List

import torch


def prepare_optimizer(optim: torch.optim.Optimizer, device: torch.device) -> None:
    """Make ``optim.step()`` capturable: only optimisers whose fresh state is all zeros are
    accepted (the guard zero-fills state that did not exist before the warm-up)."""
    ok = isinstance(optim, (torch.optim.Adam, torch.optim.AdamW, torch.optim.RMSprop)) or (
        isinstance(optim, torch.optim.SGD)
        and all(g["dampening"] == 0 or g["momentum"] == 0 for g in optim.param_groups))
    if not ok:
        raise ValueError("cuda_graph=True supports Adam, AdamW, RMSprop and SGD without "
                         f"dampening (their fresh state is all zeros); got {type(optim)}")
    for g in optim.param_groups:
        if "capturable" in g:
            g["capturable"] = True
    for st in optim.state.values():  # an optimiser that already stepped on the host
        if torch.is_tensor(st.get("step")) and not st["step"].is_cuda:
            st["step"] = st["step"].to(device)


def lr_key(optim: torch.optim.Optimizer) -> tuple:
    """Learning rates are baked into a captured ``step()``: graphs are keyed by them."""
    return tuple(float(g["lr"]) for g in optim.param_groups)


class StateGuard:
    def __init__(self, tensors: Iterable[torch.Tensor], optim: torch.optim.Optimizer) -> None:
        self.live: List[torch.Tensor] = list(tensors)
        self.saved = [t.detach().clone() for t in self.live]
        self.optim = optim
        self.saved_opt = {q: {k: v.detach().clone() for k, v in st.items() if torch.is_tensor(v)}
                          for q, st in optim.state.items()}

    def restore(self) -> None:
        with torch.no_grad():
            for t, v in zip(self.live, self.saved):
                t.copy_(v)
            for q, st in self.optim.state.items():
                for k, v in st.items():
                    if torch.is_tensor(v):
                        old = self.saved_opt.get(q, {}).get(k)
                        if old is not None:
                            v.copy_(old)
                        else:
                            v.zero_()


@contextmanager
def no_distribution_validation() -> Iterator[None]:
    """Argument validation of ``torch.distributions`` reads a device flag on the host
    (``if not valid.all()``), which a stream capture cannot contain."""
    validate = torch.distributions.Distribution._validate_args
    torch.distributions.Distribution.set_default_validate_args(False)
    try:
        yield
    finally:
        torch.distributions.Distribution.set_default_validate_args(validate)


def capture(body: Callable[[], Any], guard_tensors: Callable[[], Iterable[torch.Tensor]],
            optim: torch.optim.Optimizer, device: torch.device,
            before_each: Callable[[], None] = lambda: None, warmup: int = 3
            ) -> torch.cuda.CUDAGraph:
    """Warm ``body`` up on a side stream, restore, capture it, restore again."""
    guard = StateGuard(guard_tensors(), optim)
    side = torch.cuda.Stream(device=device)
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(warmup):
            before_each()
            body()
    torch.cuda.current_stream().wait_stream(side)
    # state created by the warm-up (fresh optimiser moments) now exists: re-guard it as zeros
    guard.restore()
    for p in (q for g in optim.param_groups for q in g["params"]):
        p.grad = None
    graph = torch.cuda.CUDAGraph()
    before_each()
    with torch.cuda.graph(graph):
        body()
    guard.restore()  # capture launches nothing; host-side resets done by before_each are undone
    return graph
