"""Replay buffer with the reference's interface (pokeai/ai/generic_move_model/replay_buffer.py:7-32).

Items keep the reference's field names.  ``state`` / ``next_state`` hold the 144-byte packed
record of the observation (a numpy void of PK_OBS_DTYPE) instead of the dense (617, A) float
array: the update kernel re-encodes on the fly, which cuts the buffer from 44 KB to 288 B per
transition.  Dense arrays are accepted too (reference checkpoints) and are
converted when the trainer ingests them.
"""
from __future__ import annotations

import random
from collections import deque
from typing import Iterable, List, NamedTuple, Optional

import numpy as np


class ReplayBufferItem(NamedTuple):
    state: np.ndarray                      # packed record (or dense (C, A) float32 from a reference checkpoint)
    action_mask: np.ndarray                # (A,) int32, 1 = legal
    action: int
    next_state: Optional[np.ndarray]       # None at the end of an episode
    next_action_mask: Optional[np.ndarray]
    reward: float


class ReplayBuffer:
    def __init__(self, size: Optional[int]):
        self.buffer = deque(maxlen=size)

    def __len__(self):
        return len(self.buffer)

    def append(self, item: ReplayBufferItem):
        self.buffer.append(item)

    def extend(self, items: Iterable[ReplayBufferItem]):
        self.buffer.extend(items)

    def sample(self, size: int) -> List[ReplayBufferItem]:
        if len(self) < size:
            raise IndexError
        return random.sample(self.buffer, size)
