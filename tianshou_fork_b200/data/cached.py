"""``CachedReplayBuffer`` on device-resident storage -- drop-in for
``tianshou.data.CachedReplayBuffer`` (tianshou/data/buffer/cached.py:8-96).

Layout ``| main_buffer | cached_buffers[0] | ... | cached_buffers[n - 1] |`` in one contiguous
device allocation; a step is written into the cached rings, and when an episode ends its
transitions are moved into the main ring (``main_buffer.update``, base.py:170-193) and the
cached ring is reset.  ``prev`` / ``next`` / ``unfinished_index`` / ``sample_indices`` are the
manager's kernels over unequal sub-buffer sizes (index goldens: test/base/test_buffer.py:904-1031).
"""
from __future__ import annotations

from typing import Any, List, Optional, Tuple, Union

import numpy as np

from .batch import Batch
from .buffer import ReplayBuffer, _host
from .manager import ReplayBufferManager


class CachedReplayBuffer(ReplayBufferManager):
    def __init__(self, main_buffer: ReplayBuffer, cached_buffer_num: int,
                 max_episode_length: int) -> None:
        assert cached_buffer_num > 0 and max_episode_length > 0
        assert type(main_buffer) == ReplayBuffer  # noqa: E721 (cached.py:41)
        kwargs = dict(main_buffer.options)
        buffers = [main_buffer] + [
            ReplayBuffer(max_episode_length, device=main_buffer.device,
                         rew_dtype=main_buffer.rew_dtype, **kwargs)
            for _ in range(cached_buffer_num)
        ]
        super().__init__(buffer_list=buffers)
        self.main_buffer = self.buffers[0]
        self.cached_buffers = self.buffers[1:]
        self.cached_buffer_num = cached_buffer_num

    def __setstate__(self, state: Any) -> None:
        super().__setstate__(state)
        self.__dict__["main_buffer"] = self.buffers[0]
        self.__dict__["cached_buffers"] = self.buffers[1:]

    def add(self, batch: Batch, buffer_ids: Optional[Union[np.ndarray, List[int]]] = None
            ) -> Tuple[np.ndarray, np.ndarray, np.ndarray, np.ndarray]:
        """Add a step into ``cached_buffers[buffer_ids]``; finished episodes move to the main
        buffer.  Returns ``(current_index, episode_reward, episode_length,
        episode_start_index)`` where the entries of finished episodes refer to the main buffer
        (-1 when nothing could be moved), cached.py:52-96."""
        if buffer_ids is None:
            buf_arr = np.arange(1, 1 + self.cached_buffer_num)
        else:
            buf_arr = np.asarray(buffer_ids) + 1
        term, trunc = _host(batch.terminated), _host(batch.truncated)
        done = np.logical_or(term, trunc).astype(bool).reshape(-1)
        ptr, ep_rew, ep_len, ep_idx = super().add(batch, buffer_ids=buf_arr)
        updated_ptr, updated_ep_idx = [], []
        for buffer_idx in buf_arr[done]:
            index = self.main_buffer.update(self.buffers[buffer_idx])
            if len(index) == 0:  # unsuccessful move, replace with -1
                index = [-1]
            updated_ep_idx.append(index[0])
            updated_ptr.append(index[-1])
            self.buffers[buffer_idx].reset()
            self._ep_rew[buffer_idx], self._ep_len[buffer_idx] = 0.0, 0
            self._ep_idx[buffer_idx] = 0
            self._lengths[0] = len(self.main_buffer)
            self._lengths[buffer_idx] = 0
            self.last_index[0] = index[-1]
            self.last_index[buffer_idx] = self._offset[buffer_idx]
        ptr[done] = updated_ptr
        ep_idx[done] = updated_ep_idx
        self._version += 1
        return ptr, ep_rew, ep_len, ep_idx
