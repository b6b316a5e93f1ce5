"""B200-native (sm_100a) return / replay hot path with the Tianshou API.

Drop-in surface (reference: aai-institute/tianshou-fork, Tianshou 0.5.1):

* ``BasePolicy.compute_episodic_return`` / ``compute_nstep_return`` / ``value_mask``
* ``ReplayBuffer`` / ``ReplayBufferManager`` / ``VectorReplayBuffer``: ``sample`` / ``prev`` /
  ``next`` / ``unfinished_index`` / ``add`` / ``__getitem__``
* ``PrioritizedReplayBuffer`` / ``PrioritizedVectorReplayBuffer``: ``sample`` / ``update_weight``
* ``SegmentTree``: ``reduce`` / ``get_prefix_sum_idx`` / ``__setitem__``
* ``PPOPolicy.learn`` with the fused advantage-norm + clip-surrogate loss

All arithmetic runs in hand-written CUDA kernels (``csrc/*.cu``) behind the C-ABI declared in
``include/ts_b200.h``; there is no CPU or eager-PyTorch fallback.
"""
from . import _lib
from .data import (Batch, to_numpy, to_torch, to_torch_as,
                   PrioritizedReplayBuffer, PrioritizedReplayBufferManager,
                   PrioritizedVectorReplayBuffer, ReplayBuffer, ReplayBufferManager,
                   SegmentTree, ShardedPrioritizedVectorReplayBuffer, VectorReplayBuffer)
from .policy import A2CPolicy, BasePolicy, DQNPolicy, PGPolicy, PPOPolicy, ppo_loss

__version__ = "0.1.0"

__all__ = [
    "Batch", "to_numpy", "to_torch", "to_torch_as", "ReplayBuffer", "ReplayBufferManager", "VectorReplayBuffer",
    "PrioritizedReplayBuffer", "PrioritizedReplayBufferManager",
    "PrioritizedVectorReplayBuffer", "ShardedPrioritizedVectorReplayBuffer", "SegmentTree",
    "BasePolicy", "PGPolicy", "A2CPolicy", "PPOPolicy", "DQNPolicy", "ppo_loss",
]
