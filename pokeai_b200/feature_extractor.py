"""FeatureExtractor: the reference's observation -> (features, mask) API
(pokeai/ai/generic_move_model/feature_extractor.py:9-52) on the B200 encode kernel.

Single observations go through the same kernel as batches (a batch of 1).  The two
sub-extractors the reference exposes as attributes (``state_feature_extractor``,
``choice_to_vec``; used by party_feature/party_evaluator_quick.py:48,53) are kept as
views of the same kernel output.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple, Union

import numpy as np

from .packed import PK_OBS_DTYPE, pack_into
from .tables import N_STATE, Tables, load_tables

ObsBatch = Union[np.ndarray, Sequence]


class _StateView:
    """StateFeatureExtractor surface (state_feature_extractor.py:36-111) backed by the encode kernel."""

    def __init__(self, owner: "FeatureExtractor"):
        self._o = owner
        self.party_size = owner.party_size

    def get_dims(self) -> int:
        return N_STATE

    def get_dim_meanings(self) -> List[str]:
        return self._o.tables.state_dim_meanings()

    def transform(self, obs) -> np.ndarray:
        feat, _ = self._o.transform(obs)
        return np.ascontiguousarray(feat[:N_STATE, 0])  # the state block is the same in every column


class _ChoiceView:
    """ChoiceToVec surface (choice_to_vec.py:12-61) backed by the encode kernel."""

    def __init__(self, owner: "FeatureExtractor"):
        self._o = owner

    def get_dims(self) -> int:
        return self._o.tables.n_choice

    def get_dim_meanings(self) -> List[str]:
        return self._o.tables.choice_dim_meanings()

    def transform(self, obs) -> np.ndarray:
        feat, _ = self._o.transform(obs)
        return np.ascontiguousarray(feat[N_STATE:, :len(obs.possible_actions)])


class FeatureExtractor:
    def __init__(self, party_size: int, device=None, tables: Optional[Tables] = None):
        self.party_size = party_size
        self.tables = tables or load_tables()
        self.device = device
        self.state_feature_extractor = _StateView(self)
        self.choice_to_vec = _ChoiceView(self)
        self._session = None

    # -- shapes (feature_extractor.py:15-38) ---------------------------------------------------------
    @property
    def input_shape(self) -> Tuple[int, int]:
        return (N_STATE + self.tables.n_choice), self.output_dim

    def get_dim_meanings(self) -> List[str]:
        return ["state/" + m for m in self.tables.state_dim_meanings()] + \
               ["choice/" + m for m in self.tables.choice_dim_meanings()]

    @property
    def output_dim(self) -> int:
        return 4 + self.party_size - 1

    # -- the native session (lazily created, grown on demand) ------------------------------------------
    def session(self, min_batch: int = 1):
        from .session import DecisionSession
        import torch
        if self._session is None or self._session.max_batch < min_batch:
            if not torch.cuda.is_available():
                raise RuntimeError("pokeai_b200 needs a CUDA device: there is no CPU feature extractor here "
                                   "(the CPU restatement lives in oracle/ and is test infrastructure)")
            dev = self.device if self.device is not None else torch.cuda.current_device()
            cap = 64
            while cap < min_batch:
                cap *= 2
            old = self._session
            self._session = DecisionSession(cap, self.output_dim, dev, materialise_features=True)
            if old is not None:
                old.close()
        return self._session

    # -- packing ----------------------------------------------------------------------------------------
    def pack(self, obs_list: ObsBatch, out: Optional[np.ndarray] = None) -> np.ndarray:
        """Observation objects (or already packed records) -> (B,) PK_OBS_DTYPE records."""
        if isinstance(obs_list, np.ndarray) and obs_list.dtype == PK_OBS_DTYPE:
            if out is None:
                return obs_list
            n = len(obs_list)
            # byte copy: numpy assigns structured records field by field, which is ~20x slower
            out[:n].view(np.uint8)[...] = np.ascontiguousarray(obs_list).view(np.uint8)
            return out[:n]
        n = len(obs_list)
        rec = out[:n] if out is not None else np.zeros(n, dtype=PK_OBS_DTYPE)
        if out is not None:
            rec[...] = np.zeros((), dtype=PK_OBS_DTYPE)
        for i, obs in enumerate(obs_list):
            pack_into(rec[i], obs, self.tables, self.output_dim)
        return rec

    # -- transform (feature_extractor.py:40-52) ----------------------------------------------------------
    def transform(self, obs) -> Tuple[np.ndarray, np.ndarray]:
        feat, mask = self.transform_batch([obs])
        return feat[0], mask[0]

    def transform_batch(self, obs_list: ObsBatch) -> Tuple[np.ndarray, np.ndarray]:
        """-> feat (B, C, A) float32, mask (B, A) int32, as host arrays."""
        n = len(obs_list)
        s = self.session(n)
        self.pack(obs_list, out=s.packed)
        return s.encode(n)
