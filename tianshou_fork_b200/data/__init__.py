"""Device-resident data structures (mirror of ``tianshou.data`` for the hot path)."""
from .batch import Batch
from .segtree import SegmentTree
from .buffer import ReplayBuffer
from .manager import ReplayBufferManager, VectorReplayBuffer
from .prio import (PrioritizedReplayBuffer, PrioritizedReplayBufferManager,
                   PrioritizedVectorReplayBuffer)
from .sharded import ShardedPrioritizedVectorReplayBuffer

__all__ = [
    "Batch", "SegmentTree", "ReplayBuffer", "ReplayBufferManager", "VectorReplayBuffer",
    "PrioritizedReplayBuffer", "PrioritizedReplayBufferManager",
    "PrioritizedVectorReplayBuffer", "ShardedPrioritizedVectorReplayBuffer",
]
