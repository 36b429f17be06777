"""Simulator request -> packed record, without PossibleAction strings in between.

The reference turns every `|request|` JSON into a list of PossibleAction named tuples full of strings
(get_possible_actions, pokeai/ai/common.py:143-174), wraps them in an RLPolicyObservation
(rl_policy.py:95-98) and lets ChoiceToVec look every string up again (choice_to_vec.py:48-60).  Here the
request is walked once and the integer indices go straight into the 144-byte record the kernels read
(`pack_request`); the simulator keys ("move 2", "switch 5") are returned beside it so the chosen slot can be
mapped back.  `PackedRLPolicy` is RLPolicy (rl_policy.py:16-125) on that path, and
`PackedRLPolicy.choice_batch` is the lock-step form: one `act_batch` launch for every battle that is waiting
for a decision.

Same results as `pack_observation(RLPolicyObservation(bs, request, get_possible_actions(bs, request)))`
(tests/test_protocol.py, against vectors produced by the reference's own get_possible_actions), same
exceptions: KeyError on unknown names or a request without 'active', ValueError when the actions do not fit.
"""
from __future__ import annotations

import struct
from typing import List, Optional, Sequence, Tuple

import numpy as np

from .observation import RANK_KEYS
from .packed import FLAG_FORCE, FLAG_SWITCH, PK_OBS_BYTES, PK_OBS_DTYPE, _status_code, pack_state_into
from .tables import MAX_ACTIONS, MAX_MOVES, STATUS_CODE, STATUS_FNT, WEATHER_CODE, Tables, load_tables


def _active_move_id(move_id: str, name: str) -> str:
    # common.py:17-21: "Hidden Power Dark 70" -> hiddenpowerdark
    if move_id == "hiddenpower":
        return "hiddenpower" + name.split(" ")[2].lower()
    return move_id


def _side_move_id(move_id: str) -> str:
    # common.py:24-29: return102 -> return, frustration1 -> frustration
    if move_id.startswith("return"):
        return "return"
    if move_id.startswith("frustration"):
        return "frustration"
    return move_id


def get_possible_actions(battle_status, request: dict, tables: Optional[Tables] = None) -> list:
    """The reference's function of the same name (common.py:143-174), for callers that want the PossibleAction list
    itself (logging, `RLPolicyObservation`): same order, filtering and renames as `pack_actions_into` below.
    `battle_status` is accepted and not read, as in the reference."""
    from .observation import PossibleAction
    tables = tables or load_tables()
    force_switch = bool(request.get("forceSwitch"))
    if not force_switch:
        active = request["active"]
        trapped = active[0].get("trapped")
    else:
        active, trapped = None, False
    out = []
    for poke_idx, member in enumerate(request["side"]["pokemon"]):
        name = member["details"].split(",")[0]
        poke_id = tables.species_ids[tables.species_name_to_idx[name]]  # dex.get_pokedex_by_name(name)['id']
        all_moves = [_side_move_id(m) for m in member["moves"]]
        if member["active"]:
            if not force_switch:
                for move_idx, move in enumerate(active[0]["moves"]):
                    if not move.get("disabled"):
                        out.append(PossibleAction(f"move {move_idx + 1}", poke_id,
                                                  _active_move_id(move["id"], move["move"]), False, False, all_moves,
                                                  member["item"]))
        else:
            if member["condition"].endswith(" fnt") or trapped:
                continue
            out.append(PossibleAction(f"switch {poke_idx + 1}", poke_id, None, True, force_switch, all_moves,
                                      member["item"]))
    return out


def pack_actions_into(rec, request: dict, tables: Tables, max_actions: int = MAX_ACTIONS) -> List[str]:
    """Fill n_act / act[] of one record from a request; returns the simulator key of every slot.
    Order and filtering are get_possible_actions' (common.py:143-174): party order; the active member gives one
    slot per enabled move (none on a forced switch); a benched member gives a switch slot unless it has fainted
    or the active one is trapped."""
    force_switch = bool(request.get("forceSwitch"))
    if not force_switch:
        active = request["active"]
        trapped = active[0].get("trapped")
    else:
        active = None
        trapped = False
    keys: List[str] = []
    acts = rec["act"]
    sp_idx, mv_idx, it_idx = tables.species_name_to_idx, tables.move_to_idx, tables.item_to_idx

    def slot():
        if len(keys) >= max_actions:
            raise ValueError(f"more than {max_actions} possible actions")  # feature_extractor.py:49 broadcast error
        return acts[len(keys)]

    for poke_idx, member in enumerate(request["side"]["pokemon"]):
        species = sp_idx[member["details"].split(",")[0]]  # "Kangaskhan, L50, F" (common.py:146-147)
        if member["active"]:
            if force_switch:
                continue
            item = it_idx[member["item"]]
            for move_idx, move in enumerate(active[0]["moves"]):
                if move.get("disabled"):
                    continue
                ar = slot()
                ar["flags"] = 0
                ar["species"] = species
                ar["item"] = item
                ar["n_moves"] = 1
                ar["move"][0] = mv_idx[_active_move_id(move["id"], move["move"])]
                keys.append(f"move {move_idx + 1}")
        else:
            if member["condition"].endswith(" fnt") or trapped:
                continue
            moves = member["moves"]
            if len(moves) > MAX_MOVES:
                raise ValueError("more than 4 moves on one action")
            ar = slot()
            ar["flags"] = FLAG_SWITCH | (FLAG_FORCE if force_switch else 0)
            ar["species"] = species
            ar["item"] = it_idx[member["item"]]
            ar["n_moves"] = len(moves)
            for j, m in enumerate(moves):
                ar["move"][j] = mv_idx[_side_move_id(m)]
            keys.append(f"switch {poke_idx + 1}")
    rec["n_act"] = len(keys)
    return keys


def pack_request(battle_status, request: dict, tables: Optional[Tables] = None,
                 max_actions: int = MAX_ACTIONS) -> Tuple[np.ndarray, List[str]]:
    """(BattleStatus, request) -> ((1,) PK_OBS_DTYPE record, simulator keys)."""
    rec, keys = pack_requests([(battle_status, request)], tables, max_actions)
    return rec, keys[0]


def pack_requests(pairs: Sequence[Tuple[object, dict]], tables: Optional[Tables] = None,
                  max_actions: int = MAX_ACTIONS) -> Tuple[np.ndarray, List[List[str]]]:
    tables = tables or load_tables()
    out = np.zeros(len(pairs), dtype=PK_OBS_DTYPE)
    keys = []
    for i, (bs, request) in enumerate(pairs):
        pack_state_into(out[i], bs, tables)
        keys.append(pack_actions_into(out[i], request, tables, max_actions))
    return out, keys


_HEADER = struct.Struct("<2B2B2H2H2B14b4B")   # offsets 0..31 of pk_obs (include/pokeai_b200.h)
_ACTION = struct.Struct("<4B4H")               # one pk_action
assert _HEADER.size == 32 and _ACTION.size == 12
_ZERO_RECORD = bytes(PK_OBS_BYTES)


def _load_c_packer():
    """The C form of the function below (csrc/pkpack.c, built into _lib/_pkpack.so by pokeai_b200.build): same bytes,
    keys and exception types, ~6x faster.  None when it is not built or PK_PY_PACKER=1."""
    import os
    if os.environ.get("PK_PY_PACKER", "0") not in ("", "0"):
        return None
    try:
        import importlib.machinery
        import importlib.util
        from .build import PACK_PATH
        if not PACK_PATH.exists():
            return None
        loader = importlib.machinery.ExtensionFileLoader("_pkpack", str(PACK_PATH))
        mod = importlib.util.module_from_spec(importlib.util.spec_from_loader("_pkpack", loader))
        loader.exec_module(mod)
        return mod.pack_request_into
    except Exception:  # wrong interpreter version, missing file: the Python implementation is the same function
        return None


_C_PACK = _load_c_packer()


def pack_request_into(buf, offset: int, battle_status, request: dict, tables: Tables,
                      max_actions: int = MAX_ACTIONS) -> List[str]:
    """`pack_request` for a driver's inner loop: the 144-byte record is written at `buf[offset:]` (any writable buffer,
    e.g. the bytes of a `pinned_records` array) and the simulator keys are returned.  Runs in C when the extension is
    built (`_load_c_packer`), else `pack_request_into_py`."""
    if _C_PACK is not None:
        return _C_PACK(buf, offset, battle_status, request, tables.species_name_to_idx, tables.move_to_idx,
                       tables.item_to_idx, max_actions, STATUS_CODE, STATUS_FNT, WEATHER_CODE)
    return pack_request_into_py(buf, offset, battle_status, request, tables, max_actions)


def pack_request_into_py(buf, offset: int, battle_status, request: dict, tables: Tables,
                         max_actions: int = MAX_ACTIONS) -> List[str]:
    """`pack_request` for a driver's inner loop: writes the 144-byte record at `buf[offset:]` (any writable buffer,
    e.g. the bytes of a `pinned_records` array) with a dozen `struct.pack_into` calls instead of ~100 numpy field
    assignments (5 us instead of 34 us per request), and returns the simulator keys.  Same bytes, same order, same
    exceptions as `pack_state_into` + `pack_actions_into` (tests/test_protocol.py)."""
    bs = battle_status
    sf, so = bs.side_statuses[bs.side_friend], bs.side_statuses[bs.side_opponent]
    af, ao = sf.active, so.active
    assert af is not None and ao is not None  # state_feature_extractor.py:133,148,161,177
    if sf.total_pokes == 0 or so.total_pokes == 0 or af.hp_max == 0 or ao.hp_max == 0:
        raise ZeroDivisionError("division by zero")  # the reference divides in Python (:122,:150)
    rf, ro = af.ranks, ao.ranks
    force_switch = bool(request.get("forceSwitch"))
    if not force_switch:
        active = request["active"]
        trapped = active[0].get("trapped")
    else:
        active, trapped = None, False
    sp_idx, mv_idx, it_idx = tables.species_name_to_idx, tables.move_to_idx, tables.item_to_idx
    opp = sp_idx[ao.species]  # KeyError like dex.get_pokedex_by_name
    buf[offset:offset + PK_OBS_BYTES] = _ZERO_RECORD
    keys: List[str] = []
    pos = offset + 32
    for poke_idx, member in enumerate(request["side"]["pokemon"]):
        species = sp_idx[member["details"].split(",")[0]]
        if member["active"]:
            if force_switch:
                continue
            item = it_idx[member["item"]]
            for move_idx, move in enumerate(active[0]["moves"]):
                if move.get("disabled"):
                    continue
                if len(keys) >= max_actions:
                    raise ValueError(f"more than {max_actions} possible actions")
                _ACTION.pack_into(buf, pos, 0, species, item, 1, mv_idx[_active_move_id(move["id"], move["move"])], 0, 0, 0)
                pos += 12
                keys.append(f"move {move_idx + 1}")
        else:
            if member["condition"].endswith(" fnt") or trapped:
                continue
            moves = member["moves"]
            if len(moves) > MAX_MOVES:
                raise ValueError("more than 4 moves on one action")
            if len(keys) >= max_actions:
                raise ValueError(f"more than {max_actions} possible actions")
            m = [mv_idx[_side_move_id(x)] for x in moves]
            m += [0] * (MAX_MOVES - len(m))
            _ACTION.pack_into(buf, pos, FLAG_SWITCH | (FLAG_FORCE if force_switch else 0), species, it_idx[member["item"]],
                              len(moves), *m)
            pos += 12
            keys.append(f"switch {poke_idx + 1}")
    _HEADER.pack_into(buf, offset, sf.remaining_pokes, so.remaining_pokes, sf.total_pokes, so.total_pokes,
                      af.hp_current, ao.hp_current, af.hp_max, ao.hp_max, _status_code(af.status), _status_code(ao.status),
                      *[rf[k] for k in RANK_KEYS], *[ro[k] for k in RANK_KEYS], opp, WEATHER_CODE.get(bs.weather, 0),
                      len(keys), 0)
    return keys


class PackedRLPolicy:
    """RLPolicy (rl_policy.py:16-125) with the request packed directly.  `agent` is an AgentVal / AgentTrain of
    this package; `surrogate_reward_config` has hp_ratio / alive_ratio / only_opponent / offset_at_end
    (surrogate_reward_config.py:4-15)."""

    def __init__(self, agent, surrogate_reward_config, tables: Optional[Tables] = None):
        self.agent = agent
        self.surrogate_reward_config = surrogate_reward_config
        self.last_reward_potential: Optional[float] = None
        self.tables = tables or load_tables()
        self._max_actions = agent._feature_extractor.output_dim

    def game_start(self) -> None:
        self.last_reward_potential = None

    def _calc_reward_potential(self, battle_status) -> float:
        c = self.surrogate_reward_config
        if c.hp_ratio == 0.0 and c.alive_ratio == 0.0:
            # every term off (the default of surrogate_reward_config.py:4-15, and what rating battles use): both sides'
            # sums are x * 0.0 + y * 0.0 = 0.0 for any state, so the potential is 0.0 (-0.0 with only_opponent) and the
            # walk over the reserve members is skipped; the state checks of get_mean_hp_ratio stay
            sides = battle_status.side_statuses
            f, o = sides[battle_status.side_friend], sides[battle_status.side_opponent]
            assert f.total_pokes > 0 and o.total_pokes > 0
            assert f.active is not None and o.active is not None
            return -0.0 if c.only_opponent else 0.0
        sps = []
        for side in (battle_status.side_friend, battle_status.side_opponent):  # rl_policy.py:66-70
            ss = battle_status.side_statuses[side]
            sps.append(ss.get_mean_hp_ratio() * c.hp_ratio + ss.get_alive_ratio() * c.alive_ratio)
        return -sps[1] if c.only_opponent else sps[0] - sps[1]

    def _surrogate(self, potential: float) -> float:
        return potential - self.last_reward_potential if self.last_reward_potential is not None else 0.0

    def _choice_by_model(self, battle_status, request: dict) -> str:
        potential = self._calc_reward_potential(battle_status)
        rec, keys = pack_request(battle_status, request, self.tables, self._max_actions)
        if len(keys) == 1:  # rl_policy.py:88-92: a single choice never reaches the model
            return keys[0]
        action = self.agent.act(rec, self._surrogate(potential))
        self.last_reward_potential = potential
        return keys[action]

    choice_turn_start = _choice_by_model
    choice_force_switch = _choice_by_model

    def game_end(self, reward: float) -> None:
        if self.surrogate_reward_config.offset_at_end:
            reward = reward - self.last_reward_potential
        self.agent.stop_episode(reward)

    @staticmethod
    def choice_batch(policies: Sequence["PackedRLPolicy"], statuses: Sequence, requests: Sequence[dict]) -> List[str]:
        """Greedy decisions of many battles that share ONE validation agent, in one launch.  Equivalent to calling
        `choice_turn_start` / `choice_force_switch` on each policy in turn (an AgentVal ignores the reward, but the
        potentials are kept so that `game_end` sees the same state)."""
        if not policies:
            return []
        agent = policies[0].agent
        if any(p.agent is not agent for p in policies):
            raise ValueError("choice_batch needs policies that share one agent")
        pots = [p._calc_reward_potential(bs) for p, bs in zip(policies, statuses)]
        rec, keys = pack_requests(list(zip(statuses, requests)), policies[0].tables, policies[0]._max_actions)
        multi = [i for i, k in enumerate(keys) if len(k) != 1]
        out = [k[0] if len(k) == 1 else "" for k in keys]
        if multi:
            actions = agent.act_batch(rec[multi])
            for i, a in zip(multi, actions):
                out[i] = keys[i][int(a)]
                policies[i].last_reward_potential = pots[i]
        return out
