"""Replay buffer with the reference's interface (pokeai/ai/generic_move_model/replay_buffer.py:7-32),
stored as a ring of packed records.

The reference keeps a ``deque`` of items holding two dense (617, A) float32 arrays each (44 KB per
transition, 4.4 GB at the default capacity of 100 000).  Here a transition is two 144-byte packed
observation records plus action / reward / terminal flag: the update kernel re-encodes
on the fly, so nothing dense is ever stored.  Storage is columnar numpy so that a minibatch is a
fancy-index gather, not a Python loop.

Reference API kept: ``ReplayBufferItem`` (same field names; ``state`` / ``next_state`` are packed records),
``ReplayBuffer(size)``, ``len``, ``append``, ``extend``, ``sample`` (same use of ``random.sample``, so a
seeded run draws the same transitions as the reference), and ``.buffer`` (iterable of items, what
``Trainer.extend_replay_buffer`` reads, trainer.py:136-139).
"""
from __future__ import annotations

import random
from typing import Iterable, Iterator, List, NamedTuple, Optional

import numpy as np

from .packed import PK_OBS_DTYPE


class ReplayBufferItem(NamedTuple):
    state: np.ndarray                      # packed record (numpy void of PK_OBS_DTYPE)
    action_mask: np.ndarray                # (A,) int32, 1 = legal (prefix mask: first n_act slots)
    action: int
    next_state: Optional[np.ndarray]       # None at the end of an episode
    next_action_mask: Optional[np.ndarray]
    reward: float


def _as_record(x, mask=None) -> np.ndarray:
    """A packed record as it is; a dense (C, A) float32 feature array -- what the reference's replay buffer holds
    (replay_buffer.py:7-15, e.g. in a `resume=True` checkpoint) -- converted to a record that encodes to the same bits."""
    if isinstance(x, (np.ndarray, np.void)) and x.dtype == PK_OBS_DTYPE:
        return x
    if isinstance(x, np.ndarray) and x.ndim == 2 and mask is not None:
        from .packed import record_from_dense
        return record_from_dense(x, mask)
    raise TypeError("replay items hold packed observation records (PK_OBS_DTYPE) or dense (C, A) feature arrays with "
                    "their action mask")


def from_reference(buffer) -> "ReplayBuffer":
    """A reference ReplayBuffer (deque of dense items, replay_buffer.py:18-32) -> the packed ring."""
    if isinstance(buffer, ReplayBuffer):
        return buffer
    items = list(buffer.buffer)
    out = ReplayBuffer(getattr(buffer.buffer, "maxlen", None))
    for it in items:
        out.append(it)
    return out


class _BufferView:
    """What ``replay_buffer.buffer`` returns: a read-only sequence of ReplayBufferItem, oldest first."""

    def __init__(self, owner: "ReplayBuffer"):
        self._o = owner

    def __len__(self) -> int:
        return len(self._o)

    def __getitem__(self, j: int) -> ReplayBufferItem:
        return self._o.item(j)

    def __iter__(self) -> Iterator[ReplayBufferItem]:
        for j in range(len(self._o)):
            yield self._o.item(j)


class ReplayBuffer:
    def __init__(self, size: Optional[int], action_slots: Optional[int] = None):
        self.maxlen = size
        self._A = action_slots
        self._cap = 0
        self._start = 0
        self._len = 0
        self._state = self._next = self._action = self._reward = self._terminal = None
        self._alloc(min(size, 1024) if size is not None else 64)

    # -- storage ---------------------------------------------------------------------------------------
    def _alloc(self, cap: int) -> None:
        new = (np.zeros(cap, dtype=PK_OBS_DTYPE), np.zeros(cap, dtype=PK_OBS_DTYPE), np.zeros(cap, dtype=np.int32),
               np.zeros(cap, dtype=np.float32), np.zeros(cap, dtype=np.uint8))
        if self._len:
            idx = self._phys(np.arange(self._len))
            for dst, src in zip(new, (self._state, self._next, self._action, self._reward, self._terminal)):
                dst[:self._len] = src[idx]
        self._state, self._next, self._action, self._reward, self._terminal = new
        self._cap, self._start = cap, 0

    def _phys(self, j):
        return (self._start + j) % self._cap

    def _slot_for_append(self) -> int:
        if self._len == self._cap:
            if self.maxlen is not None and self._cap >= self.maxlen:
                p = self._start  # full ring: overwrite the oldest (deque(maxlen) semantics)
                self._start = (self._start + 1) % self._cap
                return p
            self._alloc(min(self._cap * 2, self.maxlen) if self.maxlen is not None else self._cap * 2)
        p = self._phys(self._len)
        self._len += 1
        return p

    # -- reference API -----------------------------------------------------------------------------------
    def __len__(self) -> int:
        return self._len

    def append(self, item: ReplayBufferItem) -> None:
        if self.maxlen == 0:
            return
        p = self._slot_for_append()
        self._state[p] = _as_record(item.state, item.action_mask)
        if item.next_state is None:
            self._next[p] = np.zeros((), dtype=PK_OBS_DTYPE)
            self._terminal[p] = 1
        else:
            self._next[p] = _as_record(item.next_state, item.next_action_mask)
            self._terminal[p] = 0
        self._action[p] = item.action
        self._reward[p] = item.reward
        if self._A is None and item.action_mask is not None:
            self._A = len(item.action_mask)

    def extend(self, items: Iterable[ReplayBufferItem]) -> None:
        src = items._o if isinstance(items, _BufferView) else items
        if isinstance(src, ReplayBuffer):
            n = len(src)
            if n == 0:
                return
            idx = src._phys(np.arange(n))
            self.extend_arrays(src._state[idx], src._next[idx], src._action[idx], src._reward[idx], src._terminal[idx])
            if self._A is None:
                self._A = src._A
            return
        for it in items:
            self.append(it)

    def extend_arrays(self, state, next_state, action, reward, terminal) -> None:
        """Vectorised append of n transitions (packed records + columns)."""
        n = len(state)
        if self.maxlen is not None and n > self.maxlen:
            state, next_state, action, reward, terminal = (x[n - self.maxlen:] for x in
                                                           (state, next_state, action, reward, terminal))
            n = self.maxlen
        need = self._len + n if self.maxlen is None else min(self._len + n, self.maxlen)
        if need > self._cap:
            cap = self._cap
            while cap < need:
                cap *= 2
            self._alloc(min(cap, self.maxlen) if self.maxlen is not None else cap)
        drop = max(0, self._len + n - self._cap)  # overwritten oldest entries
        self._start = (self._start + drop) % self._cap
        self._len -= drop
        p = self._phys(self._len + np.arange(n))
        self._state[p], self._next[p] = state, next_state
        self._action[p], self._reward[p], self._terminal[p] = action, reward, terminal
        self._len += n

    def sample_indices(self, size: int) -> np.ndarray:
        """Logical indices (0 = oldest) of a minibatch; random.sample over range(len) consumes the global
        ``random`` state exactly like ``random.sample(deque, size)`` in replay_buffer.py:29-32."""
        if len(self) < size:
            raise IndexError
        return np.asarray(random.sample(range(self._len), size), dtype=np.int64)

    def sample(self, size: int) -> List[ReplayBufferItem]:
        return [self.item(int(j)) for j in self.sample_indices(size)]

    def gather(self, indices: np.ndarray):
        """-> (state, next_state, action, reward, terminal) arrays of the given logical indices."""
        p = self._phys(np.asarray(indices, dtype=np.int64))
        return self._state[p], self._next[p], self._action[p], self._reward[p], self._terminal[p]

    # -- item view ------------------------------------------------------------------------------------------
    def _mask(self, rec) -> np.ndarray:
        A = self._A if self._A is not None else 9
        m = np.zeros((A,), dtype=np.int32)
        m[:int(rec["n_act"])] = 1
        return m

    def item(self, j: int) -> ReplayBufferItem:
        if not 0 <= j < self._len:
            raise IndexError(j)
        p = self._phys(j)
        st = self._state[p]
        if self._terminal[p]:
            return ReplayBufferItem(st, self._mask(st), int(self._action[p]), None, None, float(self._reward[p]))
        nx = self._next[p]
        return ReplayBufferItem(st, self._mask(st), int(self._action[p]), nx, self._mask(nx), float(self._reward[p]))

    @property
    def buffer(self) -> _BufferView:
        return _BufferView(self)
