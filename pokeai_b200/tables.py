"""Static index tables and the feature-row layout of the decision hot path.

Row order follows the reference:
  * state rows 0..49   -- pokeai/ai/state_feature_extractor.py:74-90
  * choice rows 0..566 -- pokeai/ai/generic_move_model/choice_to_vec.py:16-28
    (switch, force_switch, poke/<251>, move/<267>, item/<47>)
  * assembled tensor = state rows then choice rows
    -- pokeai/ai/generic_move_model/feature_extractor.py:16-24,46-49

The tables themselves are derived once from the reference's data/dataset JSON
by tools/make_tables.py (only list order and species->types are kept).
"""
from __future__ import annotations

import json
from functools import lru_cache
from pathlib import Path
from typing import Dict, List

import numpy as np

# state_feature_extractor.py:9-33
POKE_TYPES = ["Normal", "Fighting", "Flying", "Poison", "Ground", "Rock", "Bug", "Ghost", "Steel",
              "Fire", "Water", "Grass", "Electric", "Psychic", "Ice", "Dragon", "Dark"]
NV_CONDITIONS = ["psn", "tox", "par", "brn", "slp", "frz"]
RANKS = ["atk", "def", "spa", "spd", "spe", "accuracy", "evasion"]
WEATHERS = ["SunnyDay", "RainDance", "Sandstorm"]

# status / weather codes of the packed record (0 == "all-zero one-hot")
STATUS_CODE = {c: i + 1 for i, c in enumerate(NV_CONDITIONS)}
STATUS_FNT = 7  # 'fnt' (and anything unknown) encodes to zeros, like the reference's loop that never matches
WEATHER_CODE = {w: i + 1 for i, w in enumerate(WEATHERS)}

# state block offsets (50 rows)
ST_REMAINING = 0   # 2: friend, opponent
ST_TYPE = 2        # 17: opponent active's types
ST_HP = 19         # 2: friend, opponent
ST_NV = 21         # 12: friend 6, opponent 6
ST_RANK = 33       # 14: friend 7, opponent 7
ST_WEATHER = 47    # 3
N_STATE = 50

MAX_ACTIONS = 9    # 4 moves + 5 switches (party_size 6); the packed record always carries 9 slots
MAX_MOVES = 4


class Tables:
    """Index order of every one-hot block plus the species -> type-bitmask table."""

    def __init__(self, raw: dict):
        assert raw["poke_types"] == POKE_TYPES
        self.species_ids: List[str] = [s["id"] for s in raw["species"]]
        self.species_names: List[str] = [s["name"] for s in raw["species"]]
        self.moves: List[str] = list(raw["moves"])
        self.items: List[str] = list(raw["items"])
        self.species_id_to_idx: Dict[str, int] = {s: i for i, s in enumerate(self.species_ids)}
        # BattleStatus.active.species is the display name ('Mr. Mime'); dex.py:16 maps name -> entry
        self.species_name_to_idx: Dict[str, int] = {s: i for i, s in enumerate(self.species_names)}
        self.move_to_idx: Dict[str, int] = {m: i for i, m in enumerate(self.moves)}
        self.item_to_idx: Dict[str, int] = {m: i for i, m in enumerate(self.items)}
        self.species_types: List[List[int]] = [list(s["types"]) for s in raw["species"]]
        mask = np.zeros(len(self.species_ids), dtype=np.uint32)
        for i, ts in enumerate(self.species_types):
            for t in ts:
                mask[i] |= np.uint32(1 << t)
        self.species_type_mask = mask
        self.n_species = len(self.species_ids)
        self.n_moves = len(self.moves)
        self.n_items = len(self.items)
        assert self.n_species < 256 and self.n_items < 256 and self.n_moves < 65536
        # choice block offsets (choice_to_vec.py:17-26)
        self.CH_SWITCH = 0
        self.CH_FORCE = 1
        self.CH_SPECIES = 2
        self.CH_MOVE = self.CH_SPECIES + self.n_species
        self.CH_ITEM = self.CH_MOVE + self.n_moves
        self.n_choice = self.CH_ITEM + self.n_items
        self.n_channels = N_STATE + self.n_choice

    # -- meanings (state_feature_extractor.py:74-90, choice_to_vec.py:17-26, feature_extractor.py:26-28)
    def state_dim_meanings(self) -> List[str]:
        ms = ["remaining_count/friend", "remaining_count/opponent"]
        ms += [f"poke_type/opponent/{t}" for t in POKE_TYPES]
        ms += ["hp_ratio/friend", "hp_ratio/opponent"]
        for side in ("friend", "opponent"):
            ms += [f"nv_condition/{side}/{c}" for c in NV_CONDITIONS]
        for side in ("friend", "opponent"):
            ms += [f"rank/{side}/{r}" for r in RANKS]
        ms += [f"weather/{w}" for w in WEATHERS]
        return ms

    def choice_dim_meanings(self) -> List[str]:
        return (["switch", "force_switch"] + [f"poke/{n}" for n in self.species_ids]
                + [f"move/{n}" for n in self.moves] + [f"item/{n}" for n in self.items])


@lru_cache(maxsize=None)
def load_tables(path: str | None = None) -> Tables:
    p = Path(path) if path else Path(__file__).resolve().parent / "data" / "tables.json"
    return Tables(json.loads(p.read_text(encoding="utf-8")))
