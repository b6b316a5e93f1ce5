"""Sharded prioritized replay: one process per GPU, each holding a power-of-two-aligned
subtree of the reference's single global sum tree.

The reference keeps ONE tree over the whole capacity of a ``PrioritizedVectorReplayBuffer``
(tianshou/data/buffer/manager.py:209-214).  Here rank ``r`` of ``G`` owns sub-buffers
``[r*B/G, (r+1)*B/G)`` and the leaves ``[r*cap, (r+1)*cap)`` with ``cap`` a power of two, i.e.
exactly the subtree under node ``G + r`` of the global heap.  The ``log2(G)`` levels above are a
function of the ``G`` shard roots only, so after every tree change the ranks all-gather
``{root, max_prio, min_prio}`` (24 bytes each) and every rank rebuilds the same top levels in
the global tree's own pairwise order inside the sampling kernel
(``ts_per_sample_sharded``).  With identical uniform draws on every rank the union of the
ranks' samples is bit-identical to the reference's samples from the global tree.

Calls are collective: every rank calls ``add`` / ``sample`` / ``update_weight`` in the same
order (SPMD), as with any ``torch.distributed`` program.
"""
from __future__ import annotations

from typing import Any, Optional

import numpy as np
import torch
import torch.distributed as dist

from .. import _lib as L
from .prio import PrioritizedVectorReplayBuffer

# the ranged update kernel takes 4096 batch entries per pass (csrc/segtree.cu: kRangeUnroll *
# RANGE_NT); only larger batches gain from the sampler's owner list
OWNED_LIST_MIN_BATCH = 4096


def shard_layout(total_size: int, buffer_num: int, world_size: int, rank: int):
    """(sub-buffer size, sub-buffers per rank, leaves per rank, first global leaf of ``rank``)
    for ``PrioritizedVectorReplayBuffer(total_size, buffer_num)`` (vecbuf.py:57-63) split over
    ``world_size`` ranks.  Pure integer arithmetic, no device needed."""
    assert buffer_num > 0 and world_size > 0 and 0 <= rank < world_size
    assert buffer_num % world_size == 0, "sub-buffers must split evenly over the ranks"
    size = int(np.ceil(total_size / buffer_num))  # vecbuf.py:60
    per_rank = buffer_num // world_size
    cap = size * per_rank
    return size, per_rank, cap, rank * cap


def is_exact_layout(cap: int, world_size: int) -> bool:
    """Bit-exact equivalence with the global tree needs pow2 shards and a pow2 rank count."""
    return cap & (cap - 1) == 0 and world_size & (world_size - 1) == 0


def exchange_roots(mine3: torch.Tensor, world_size: int,
                   group: Optional[dist.ProcessGroup] = None) -> torch.Tensor:
    """All-gather of ``{root, max_prio, min_prio}``: ``[3]`` -> ``[world_size, 3]`` on the
    tensor's own device (NCCL on GPUs, gloo in the CPU tests)."""
    out = torch.empty(world_size, 3, dtype=mine3.dtype, device=mine3.device)
    if world_size == 1:
        out[0] = mine3
    else:
        dist.all_gather_into_tensor(out.view(-1), mine3.contiguous(), group=group)
    return out


class ShardedPrioritizedVectorReplayBuffer(PrioritizedVectorReplayBuffer):
    """This rank's shard of ``PrioritizedVectorReplayBuffer(total_size, buffer_num)``.

    Indices handed out and taken in are LOCAL (``0 <= idx < len(shard)``; ``-1`` marks a sample
    owned by another rank); ``to_global`` maps them into the reference's global index space.
    """

    def __init__(self, total_size: int, buffer_num: int, rank: Optional[int] = None,
                 world_size: Optional[int] = None, group: Optional[dist.ProcessGroup] = None,
                 exchange: str = "auto", **kwargs: Any) -> None:
        if world_size is None:
            world_size = dist.get_world_size(group) if dist.is_initialized() else 1
        if rank is None:
            rank = dist.get_rank(group) if dist.is_initialized() else 0
        size, per_rank, cap, offset = shard_layout(total_size, buffer_num, world_size, rank)
        assert is_exact_layout(cap, world_size), (
            "each shard must be a power-of-two-aligned subtree of the global tree "
            f"(shard capacity {cap}, {world_size} ranks): see SURVEY.md 8e")
        super().__init__(cap, per_rank, **kwargs)
        self.rank, self.world_size, self.group, self.offset = rank, world_size, group, offset
        self._mine3 = torch.zeros(3, dtype=torch.float64, device=self.device)
        self._roots3 = torch.zeros(world_size, 3, dtype=torch.float64, device=self.device)
        self._wmax = torch.zeros(1, dtype=torch.float64, device=self.device)
        # exchange over NVLink peer memory inside the kernels ("p2p"), or NCCL collectives
        # between them ("nccl"); "auto" tries p2p when there is more than one rank
        assert exchange in ("auto", "p2p", "nccl")
        self.exchange = "nccl"
        if world_size > 1 and exchange in ("auto", "p2p"):
            try:
                self._setup_p2p()
                self.exchange = "p2p"
            except Exception:
                if exchange == "p2p":
                    raise
        self.sync_roots()

    def _setup_p2p(self) -> None:
        """Slot buffer every peer can write: torch symmetric memory (cudaIpc / fabric handles
        exchanged through the process group's store), zeroed before anyone publishes."""
        import torch.distributed._symmetric_memory as symm_mem
        grp = self.group if self.group is not None else dist.group.WORLD
        self._slots = symm_mem.empty(16 * self.world_size, dtype=torch.float64,
                                     device=self.device)
        self._slots.zero_()
        self._symm = symm_mem.rendezvous(self._slots, grp)
        self._peer_ptrs = int(self._symm.buffer_ptrs_dev)
        self._p2p_state = torch.zeros(2, dtype=torch.int64, device=self.device)
        # batch positions this rank owns, handed from the sampling kernel to the update of the
        # same batch (int32[4 + capacity], see ts_per_sample_p2p_owned); grown on demand
        self._owned: Optional[torch.Tensor] = None
        self._owned_src: Optional[tuple] = None  # (index tensor handed out, its version)
        torch.cuda.synchronize(self.device)
        dist.barrier(group=self.group)

    # ---- the one exchange step of the path ---------------------------------------------------
    def sync_roots(self) -> None:
        """Publish this shard's root and priority extrema, adopt the global extrema (the
        reference has ONE ``_max_prio`` / ``_min_prio``, prio.py:35, 88-89).  Stream-ordered
        device copies + one 24-byte all-gather; no host sync."""
        if self.exchange == "p2p":
            L.check(L.lib().ts_per_publish(
                L.ptr(self.weight._value), L.ptr(self._prio), L.ptr(self._p2p_state),
                self._peer_ptrs, self.world_size, self.rank, L.stream()), "ts_per_publish")
            return  # the global extrema are adopted by the next sampling kernel
        self._mine3[0:1].copy_(self.weight._value[1:2])
        self._mine3[1:3].copy_(self._prio)
        self._roots3 = exchange_roots(self._mine3, self.world_size, self.group)
        self._prio[0:1].copy_(self._roots3[:, 1].amax(0, keepdim=True))
        self._prio[1:2].copy_(self._roots3[:, 2].amin(0, keepdim=True))

    def to_global(self, idx: torch.Tensor) -> torch.Tensor:
        return torch.where(idx >= 0, idx + self.offset, idx)

    # ---- collective versions of the PER calls ------------------------------------------------
    def add(self, batch, buffer_ids=None):
        out = super().add(batch, buffer_ids)  # init_weight(ptr) inside
        self.sync_roots()
        return out

    def add_device(self, batch, buffer_ids=None):
        out = super().add_device(batch, buffer_ids)  # init_weight(ptrs) inside
        self.sync_roots()
        return out

    def sample_with_uniforms(self, u: torch.Tensor) -> torch.Tensor:
        """``u`` must hold the SAME draws on every rank (same seed, or broadcast)."""
        assert u.is_cuda and u.dtype == torch.float64 and u.is_contiguous()
        m = u.numel()
        idx = torch.empty(m, dtype=torch.int64, device=self.device)
        w = torch.empty(m, dtype=torch.float64, device=self.device)
        t = self.weight
        if self.exchange == "p2p":
            # a batch the update kernel would scan in several passes: let the sampler list the
            # positions this rank owns (~ m / world_size of them) for it
            owned = 0
            self._owned_src = None
            if m > OWNED_LIST_MIN_BATCH:
                if self._owned is None or self._owned.numel() < 4 + m:
                    self._owned = torch.zeros(4 + m, dtype=torch.int32, device=self.device)
                owned = L.ptr(self._owned)
                self._owned_src = (idx, idx._version)
            L.check(L.lib().ts_per_sample_p2p_owned(
                L.ptr(t._value), t._bound, L.ptr(u), m, L.ptr(self._slots), self._peer_ptrs,
                L.ptr(self._p2p_state), self.world_size, self.rank, L.ptr(self._prio),
                float(self._beta), int(self._weight_norm), L.ptr(idx), L.ptr(w), owned,
                L.ptr(t._err), L.stream()), "ts_per_sample_p2p_owned")
            self._cached_weight = (idx, w)
            return idx
        L.check(L.lib().ts_per_sample_sharded(
            L.ptr(t._value), t._bound, L.ptr(u), m, L.ptr(self._roots3), self.world_size,
            self.rank, L.ptr(self._prio), float(self._beta), int(self._weight_norm), L.ptr(idx),
            L.ptr(w), L.ptr(self._wmax), L.ptr(t._err), L.stream()), "ts_per_sample_sharded")
        if self._weight_norm:
            if self.world_size > 1:
                dist.all_reduce(self._wmax, op=dist.ReduceOp.MAX, group=self.group)
            L.check(L.lib().ts_per_normalize(L.ptr(w), m, L.ptr(self._wmax), L.stream()),
                    "ts_per_normalize")
        self._cached_weight = (idx, w)
        return idx

    def update_weight(self, index, new_weight) -> None:
        """``index`` local, ``-1`` entries (other ranks' samples) are skipped by the kernel."""
        if self.exchange != "p2p":
            super().update_weight(index, new_weight)
            self.sync_roots()
            return
        idx = self.weight._dev_index(index)
        td = new_weight.detach().to(self.device) if isinstance(new_weight, torch.Tensor) \
            else torch.as_tensor(np.asarray(new_weight), device=self.device)
        if td.dtype not in (torch.float32, torch.float64):
            td = td.to(torch.float64)
        td = td.reshape(-1).contiguous()
        assert td.numel() == idx.numel()
        self._cached_weight = None
        t = self.weight
        # the sampler's owner list is valid only for the very tensor it filled, unmodified
        src = self._owned_src
        owned = L.ptr(self._owned) if (src is not None and index is src[0]
                                       and index._version == src[1]
                                       and idx.data_ptr() == index.data_ptr()) else 0
        # update + publish in one launch (the record must go out even for an empty batch)
        L.check(L.lib().ts_per_update_publish_owned(
            L.ptr(t._value), t._size, t._bound, L.ptr(idx), L.ptr(td), L.dtype_code(td),
            idx.numel(), float(self._alpha), L.ptr(self._prio), L.ptr(t._stamp),
            L.ptr(t._counter), L.ptr(t._err), L.ptr(self._p2p_state), self._peer_ptrs,
            self.world_size, self.rank, owned, L.stream()), "ts_per_update_publish_owned")
        self._prio_f32 = td.dtype == torch.float32
