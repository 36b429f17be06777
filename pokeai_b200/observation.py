"""Input record types of the decision path (the reference's observation objects).

These mirror, field for field, what the reference's encoder reads:
  * BattleStatus / SideStatus / ActivePokeStatus -- pokeai/ai/battle_status.py:41-188
  * PossibleAction                               -- pokeai/ai/common.py:7-14
  * RLPolicyObservation                          -- pokeai/ai/rl_policy_observation.py:7-10

The packer (pokeai_b200/packed.py) is duck-typed, so the reference's own objects
work unchanged; these classes exist so the path can be driven (tests, synthetic
benchmarks, fixtures) on a machine where the reference is not installed.
"""
from __future__ import annotations

from typing import Dict, List, NamedTuple, Optional, Set

RANK_KEYS = ("atk", "def", "spa", "spd", "spe", "accuracy", "evasion")


class PossibleAction(NamedTuple):
    simulator_key: str
    poke: str
    move: Optional[str]
    switch: bool
    force_switch: bool
    allMoves: List[str]
    item: str


class ActivePokeStatus:
    def __init__(self, pokemon: str, species: str, level: int, gender: str, hp_current: int, hp_max: int,
                 status: str):
        self.pokemon = pokemon
        self.species = species  # display name, e.g. 'Mr. Mime'
        self.level = level
        self.gender = gender
        self.hp_current = hp_current
        self.hp_max = hp_max
        self.status = status
        self.ranks: Dict[str, int] = {k: 0 for k in RANK_KEYS}
        self.volatile_statuses: Set[str] = set()

    @property
    def hp_ratio(self) -> float:  # battle_status.py:96
        return self.hp_current / self.hp_max


class SideStatus:
    def __init__(self):
        self.active: Optional[ActivePokeStatus] = None
        self.reserve_pokes: Dict[str, ActivePokeStatus] = {}
        self.side_statuses: Set[str] = set()
        self.total_pokes = 0
        self.remaining_pokes = 0

    # the two potentials RLPolicy's surrogate reward reads (battle_status.py:132-150)
    def get_mean_hp_ratio(self) -> float:
        assert self.total_pokes > 0
        assert self.active is not None
        # members that have not appeared yet count as full HP
        ratio_sum = self.active.hp_ratio + sum(p.hp_ratio for p in self.reserve_pokes.values()) + (
            self.total_pokes - 1 - len(self.reserve_pokes))
        return ratio_sum / self.total_pokes

    def get_alive_ratio(self) -> float:
        assert self.total_pokes > 0
        return self.remaining_pokes / self.total_pokes


class BattleStatus:
    WEATHER_NONE = "none"

    def __init__(self, side_friend: str, side_party=None):
        assert side_friend in ("p1", "p2")
        self.side_friend = side_friend
        self.side_opponent = "p2" if side_friend == "p1" else "p1"
        self.side_party = side_party
        self.turn = 0
        self.weather = BattleStatus.WEATHER_NONE
        self.side_statuses: Dict[str, SideStatus] = {"p1": SideStatus(), "p2": SideStatus()}


class RLPolicyObservation(NamedTuple):
    battle_status: BattleStatus
    request: dict
    possible_actions: List[PossibleAction]


def battle_status_from_json(d: dict) -> BattleStatus:
    """Rebuild a BattleStatus from the dict form the reference logs
    (``BattleStatus.json_dumps``, battle_status.py:181-188; format_battle_log.py:70-72)."""
    bs = BattleStatus(d["side_friend"], d.get("side_party"))
    bs.turn = d.get("turn", 0)
    bs.weather = d.get("weather", BattleStatus.WEATHER_NONE)
    for side, sd in d["side_statuses"].items():
        ss = SideStatus()
        ss.total_pokes = sd["total_pokes"]
        ss.remaining_pokes = sd["remaining_pokes"]
        ss.side_statuses = set(sd.get("side_statuses", []))

        def mk(a):
            ap = ActivePokeStatus(a["pokemon"], a["species"], a["level"], a["gender"], a["hp_current"],
                                  a["hp_max"], a["status"])
            ap.ranks = dict(a["ranks"])
            ap.volatile_statuses = set(a.get("volatile_statuses", []))
            return ap

        ss.active = mk(sd["active"]) if sd.get("active") is not None else None
        ss.reserve_pokes = {k: mk(v) for k, v in sd.get("reserve_pokes", {}).items()}
        bs.side_statuses[side] = ss
    return bs


def possible_actions_from_json(lst: List[dict]) -> List[PossibleAction]:
    return [PossibleAction(a["simulator_key"], a["poke"], a["move"], bool(a["switch"]), bool(a["force_switch"]),
                           list(a["allMoves"]), a["item"]) for a in lst]
