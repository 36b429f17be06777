"""ORACLE (test infrastructure, not product code) -- numpy restatement of the encoder
working directly on packed records, vectorised over the batch so that parity can be
checked at sizes where the object-level port (oracle/ref_port.py) is too slow.

Pinned against ref_port.transform (itself pinned against the reference's golden
vectors) by tests/test_oracle.py::test_packed_oracle_equals_port.

Restates: state_feature_extractor.py:92-197, choice_to_vec.py:40-61,
feature_extractor.py:40-52 of the reference (see ref_port.py for line-level notes).
"""
from __future__ import annotations

from typing import Optional, Tuple

import numpy as np

from pokeai_b200.packed import FLAG_FORCE, FLAG_SWITCH
from pokeai_b200.tables import (N_STATE, ST_HP, ST_NV, ST_RANK, ST_REMAINING, ST_TYPE, ST_WEATHER, Tables,
                                load_tables)


def state_block(rec: np.ndarray, tables: Tables) -> np.ndarray:
    """(B,) records -> (B, 50) float32."""
    B = rec.shape[0]
    st = np.zeros((B, N_STATE), dtype=np.float32)
    # python-double division rounded once to float32, as the reference does
    st[:, ST_REMAINING:ST_REMAINING + 2] = (rec["remaining"].astype(np.float64) / rec["total"].astype(np.float64))
    tm = tables.species_type_mask[rec["opp_species"]]
    for t in range(17):
        st[:, ST_TYPE + t] = (tm >> np.uint32(t)) & np.uint32(1)
    st[:, ST_HP:ST_HP + 2] = rec["hp_cur"].astype(np.float64) / rec["hp_max"].astype(np.float64)
    for side in range(2):
        code = rec["status"][:, side].astype(np.int64)
        ok = (code >= 1) & (code <= 6)
        rows = np.nonzero(ok)[0]
        st[rows, ST_NV + 6 * side + code[rows] - 1] = 1.0
        st[:, ST_RANK + 7 * side: ST_RANK + 7 * side + 7] = (rec["boost"][:, side, :].astype(np.float64) + 6.0) / 12.0
    w = rec["weather"].astype(np.int64)
    rows = np.nonzero((w >= 1) & (w <= 3))[0]
    st[rows, ST_WEATHER + w[rows] - 1] = 1.0
    return st


def encode(rec: np.ndarray, party_size: int, tables: Optional[Tables] = None) -> Tuple[np.ndarray, np.ndarray]:
    """(B,) records -> feat (B, C, A) float32, mask (B, A) int32."""
    tables = tables or load_tables()
    A = 4 + party_size - 1
    B = rec.shape[0]
    C = tables.n_channels
    feat = np.zeros((B, C, A), dtype=np.float32)
    feat[:, :N_STATE, :] = state_block(rec, tables)[:, :, None]
    n_act = rec["n_act"].astype(np.int64)
    assert (n_act <= A).all()
    mask = (np.arange(A)[None, :] < n_act[:, None]).astype(np.int32)
    bidx = np.arange(B)
    for c in range(A):
        legal = np.nonzero(mask[:, c])[0]
        if len(legal) == 0:
            continue
        a = rec["act"][legal, c]
        fl = a["flags"]
        feat[legal, N_STATE + tables.CH_SWITCH, c] = (fl & FLAG_SWITCH) != 0
        feat[legal, N_STATE + tables.CH_FORCE, c] = (fl & FLAG_FORCE) != 0
        feat[legal, N_STATE + tables.CH_SPECIES + a["species"].astype(np.int64), c] = 1.0
        feat[legal, N_STATE + tables.CH_ITEM + a["item"].astype(np.int64), c] = 1.0
        for m in range(4):
            has = a["n_moves"] > m
            rows = legal[has]
            feat[rows, N_STATE + tables.CH_MOVE + a["move"][has, m].astype(np.int64), c] = 1.0
    return feat, mask
