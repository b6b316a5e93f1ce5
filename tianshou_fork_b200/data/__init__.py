"""Device-resident data structures (mirror of ``tianshou.data`` for the hot path)."""
from .batch import Batch
from .segtree import SegmentTree
from .buffer import ReplayBuffer
from .manager import ReplayBufferManager, VectorReplayBuffer
from .cached import CachedReplayBuffer
from .prio import (PrioritizedReplayBuffer, PrioritizedReplayBufferManager,
                   PrioritizedVectorReplayBuffer)
from .sharded import ShardedPrioritizedVectorReplayBuffer

__all__ = [
    "Batch", "SegmentTree", "ReplayBuffer", "ReplayBufferManager", "VectorReplayBuffer",
    "CachedReplayBuffer",
    "PrioritizedReplayBuffer", "PrioritizedReplayBufferManager",
    "PrioritizedVectorReplayBuffer", "ShardedPrioritizedVectorReplayBuffer",
]
