"""Device-resident data structures (mirror of ``tianshou.data`` for the hot path)."""
from .batch import Batch
from .converter import to_numpy, to_torch, to_torch_as
from .segtree import SegmentTree
from .buffer import ReplayBuffer
from .manager import ReplayBufferManager, VectorReplayBuffer
from .prio import (PrioritizedReplayBuffer, PrioritizedReplayBufferManager,
                   PrioritizedVectorReplayBuffer)
from .sharded import ShardedPrioritizedVectorReplayBuffer

__all__ = [
    "Batch", "to_numpy", "to_torch", "to_torch_as", "SegmentTree", "ReplayBuffer", "ReplayBufferManager", "VectorReplayBuffer",
    "PrioritizedReplayBuffer", "PrioritizedReplayBufferManager",
    "PrioritizedVectorReplayBuffer", "ShardedPrioritizedVectorReplayBuffer",
]
