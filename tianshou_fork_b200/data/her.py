"""Hindsight Experience Replay (arXiv:1707.01495, 'future' strategy, online sampling) on
device-resident storage -- drop-ins for ``tianshou.data.HERReplayBuffer``
(tianshou/data/buffer/her.py:8-197), ``HERReplayBufferManager`` (manager.py:217-256) and
``HERVectorReplayBuffer`` (vecbuf.py:71-91).

Observations are dicts with ``observation`` / ``achieved_goal`` / ``desired_goal``.  Sampling
re-writes the goals (and rewards) of the sampled transitions' episodes in the storage and
restores them before the next write or sample, exactly as the reference does, so that n-step
returns etc. see consistent data.  The episode walk is ``horizon - 1`` calls of the ``next``
kernel; the index bookkeeping (sort, unique, the two host RNG draws) stays on the host in the
reference's order, so a seeded run re-writes the same episodes with the same goals; the goal /
reward rows themselves are gathered and scattered on the device.
"""
from __future__ import annotations

from typing import Any, Callable, List, Optional, Sequence, Tuple, Union

import numpy as np
import torch

from .batch import Batch
from .buffer import ReplayBuffer
from .manager import ReplayBufferManager


class HERReplayBuffer(ReplayBuffer):
    """``compute_reward_fn(achieved_goal, desired_goal) -> rewards`` takes and returns NumPy
    arrays of shape ``(batch, ...)`` / ``(batch,)`` like the reference's (her.py:17-20)."""

    def __init__(self, size: int, compute_reward_fn: Callable[[np.ndarray, np.ndarray], np.ndarray],
                 horizon: int, future_k: float = 8.0, **kwargs: Any) -> None:
        super().__init__(size, **kwargs)
        self.horizon = horizon
        self.future_p = 1 - 1 / future_k
        self.compute_reward_fn = compute_reward_fn
        self._original_meta = Batch()
        self._altered_indices = np.array([])

    # ---- cache of the original rows (her.py:48-82) -------------------------------------------
    def _restore_cache(self) -> None:
        """Write the cached original rows back; called before writing, sampling or saving."""
        if "_altered_indices" not in self.__dict__:
            return
        if self._altered_indices.size == 0:
            return
        self._meta[self._altered_indices] = self._original_meta
        self._original_meta = Batch()
        self._altered_indices = np.array([])
        self._version += 1

    def reset(self, keep_statistics: bool = False) -> None:
        self._restore_cache()
        return super().reset(keep_statistics)

    def set_batch(self, batch: Batch) -> None:
        self._restore_cache()
        return super().set_batch(batch)

    def update(self, buffer: ReplayBuffer) -> np.ndarray:
        self._restore_cache()
        return super().update(buffer)

    def add(self, batch: Batch, buffer_ids: Optional[Union[np.ndarray, List[int]]] = None
            ) -> Tuple[np.ndarray, np.ndarray, np.ndarray, np.ndarray]:
        self._restore_cache()
        return super().add(batch, buffer_ids)

    def __getstate__(self):
        self._restore_cache()  # what save_hdf5 does in the reference (her.py:66-68)
        return super().__getstate__()

    # ---- sampling (her.py:84-96) -------------------------------------------------------------
    def sample_indices(self, batch_size: int) -> torch.Tensor:
        self._restore_cache()
        indices = super().sample_indices(batch_size=batch_size)
        self.rewrite_transitions(indices.cpu().numpy().copy())
        return indices

    def rewrite_transitions(self, indices: np.ndarray) -> None:
        """Re-write the goal of some sampled transitions' episodes (her.py:98-189)."""
        if indices.size == 0:
            return
        dev = self.device
        # sort by age: slots before the write pointer are the newest part of a wrapped ring
        indices[indices < self._index] += self.maxsize
        indices = np.sort(indices)
        indices[indices >= self.maxsize] -= self.maxsize

        steps = [indices]
        for _ in range(self.horizon - 1):
            steps.append(np.asarray(self.next(steps[-1])))
        walk = np.stack(steps)  # [horizon, bsz]: every sample's episode from its step on

        current, terminal = walk[0], walk[-1]
        episodes_len = (terminal - current + self.maxsize) % self.maxsize
        future_offset = np.random.uniform(size=len(walk[0])) * episodes_len
        future_offset = np.round(future_offset).astype(int)
        future_t = (current + future_offset) % self.maxsize

        # one column per distinct episode; the LAST sample of an episode supplies its goal
        unique_ep_open_indices = np.sort(np.unique(terminal, return_index=True)[1])
        unique_ep_indices = walk[:, unique_ep_open_indices]
        unique_ep_close_indices = np.hstack(
            [(unique_ep_open_indices - 1)[1:], len(terminal) - 1])
        her_ep_indices = np.random.choice(
            len(unique_ep_open_indices),
            size=int(len(unique_ep_open_indices) * self.future_p), replace=False)

        self._altered_indices = unique_ep_indices.copy()
        self._original_meta = self._meta[self._altered_indices].copy()

        episodes = self[unique_ep_indices]
        ep_obs, ep_rew = episodes.obs, episodes.rew
        her = torch.from_numpy(np.asarray(her_ep_indices, dtype=np.int64)).to(dev)
        if self._save_obs_next:
            ep_obs_next = episodes.obs_next
            future_obs = self[future_t[unique_ep_close_indices]].obs_next
        else:
            future_obs = self[np.asarray(self.next(future_t[unique_ep_close_indices]))].obs

        ep_obs.desired_goal[:, her] = future_obs.achieved_goal[None, her]
        if self._save_obs_next:
            ep_obs_next.desired_goal[:, her] = future_obs.achieved_goal[None, her]
            ep_rew[:, her] = self._compute_reward(ep_obs_next)[:, her].to(ep_rew.dtype)
        else:
            tmp_ep_obs_next = self[np.asarray(self.next(unique_ep_indices))].obs
            ep_rew[:, her] = self._compute_reward(tmp_ep_obs_next)[:, her].to(ep_rew.dtype)

        assert tuple(ep_obs.desired_goal.shape[:2]) == unique_ep_indices.shape
        assert tuple(ep_obs.achieved_goal.shape[:2]) == unique_ep_indices.shape
        assert tuple(ep_rew.shape) == unique_ep_indices.shape

        where = torch.from_numpy(unique_ep_indices).to(dev)
        self._meta.obs[where] = ep_obs
        if self._save_obs_next:
            self._meta.obs_next[where] = ep_obs_next
        # the reference stores ep_rew.astype(np.float32) into its float storage (her.py:189)
        self._meta.rew[where] = ep_rew.to(torch.float32).to(self._meta.rew.dtype)
        self._version += 1

    def _compute_reward(self, obs: Batch, lead_dims: int = 2) -> torch.Tensor:
        lead_shape = tuple(obs.observation.shape[:lead_dims])
        g = obs.desired_goal.reshape(-1, *obs.desired_goal.shape[lead_dims:])
        ag = obs.achieved_goal.reshape(-1, *obs.achieved_goal.shape[lead_dims:])
        rewards = np.asarray(self.compute_reward_fn(ag.cpu().numpy(), g.cpu().numpy()))
        rewards = rewards.reshape(*lead_shape, *rewards.shape[1:])
        return torch.from_numpy(np.ascontiguousarray(rewards)).to(self.device)


class HERReplayBufferManager(ReplayBufferManager):
    """A list of HERReplayBuffers over one contiguous storage (manager.py:217-256)."""

    def __init__(self, buffer_list: Sequence[HERReplayBuffer]) -> None:
        super().__init__(list(buffer_list))

    def _restore_cache(self) -> None:
        for buf in self.buffers:
            buf._restore_cache()

    def set_batch(self, batch: Batch) -> None:
        self._restore_cache()
        return super().set_batch(batch)

    def add(self, batch: Batch, buffer_ids: Optional[Union[np.ndarray, List[int]]] = None
            ) -> Tuple[np.ndarray, np.ndarray, np.ndarray, np.ndarray]:
        self._restore_cache()
        return super().add(batch, buffer_ids)

    def __getstate__(self):
        self._restore_cache()
        return super().__getstate__()

    def sample_indices(self, batch_size: int) -> torch.Tensor:
        """manager.py:163-192 with the children's own ``sample_indices`` (each re-writes the
        episodes it sampled); host RNG calls in the reference's order."""
        if batch_size < 0:
            return torch.empty(0, dtype=torch.int64, device=self.device)
        if self._sample_avail and self.stack_num > 1:
            all_indices = torch.cat([buf.sample_indices(0) + int(offset)
                                     for offset, buf in zip(self._offset, self.buffers)])
            if batch_size == 0:
                return all_indices
            pick = np.random.choice(all_indices.numel(), batch_size)
            return all_indices[torch.from_numpy(pick).to(self.device)]
        if batch_size == 0:
            sample_num = np.zeros(self.buffer_num, int)
        else:
            buffer_idx = np.random.choice(
                self.buffer_num, batch_size, p=self._lengths / self._lengths.sum())
            sample_num = np.bincount(buffer_idx, minlength=self.buffer_num)
            sample_num[sample_num == 0] = -1  # avoid batch_size == 0: no draw, no rewrite
        parts = [buf.sample_indices(int(bsz)) + int(offset)
                 for offset, buf, bsz in zip(self._offset, self.buffers, sample_num)]
        self._version += 1
        return torch.cat(parts) if parts else torch.empty(0, dtype=torch.int64,
                                                          device=self.device)


class HERVectorReplayBuffer(HERReplayBufferManager):
    """``buffer_num`` equal HERReplayBuffers (vecbuf.py:71-91)."""

    def __init__(self, total_size: int, buffer_num: int, **kwargs: Any) -> None:
        assert buffer_num > 0
        size = int(np.ceil(total_size / buffer_num))
        buffer_list = [HERReplayBuffer(size, **kwargs) for _ in range(buffer_num)]
        super().__init__(buffer_list)
