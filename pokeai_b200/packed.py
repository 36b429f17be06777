"""Packed observation record (144 B) and its host-side packer.

The reference hands the encoder Python objects full of strings
(RLPolicyObservation: pokeai/ai/rl_policy_observation.py:7-10; the fields the
encoder reads are listed in SURVEY.md §8 a1).  The B200 path consumes a fixed
little-endian record instead so that one battle is nine 16-byte loads:

    offset size  field
    0      2     remaining[2]      friend, opponent   (SideStatus.remaining_pokes)
    2      2     total[2]                              (SideStatus.total_pokes)
    4      4     hp_cur[2]  u16                        (ActivePokeStatus.hp_current)
    8      4     hp_max[2]  u16                        (ActivePokeStatus.hp_max)
    12     2     status[2]         0 none, 1..6 psn,tox,par,brn,slp,frz, 7 fnt/other
    14     14    boost[2][7] i8    atk,def,spa,spd,spe,accuracy,evasion
    28     1     opp_species       index into all_pokemons (types via device table)
    29     1     weather           0 none, 1 SunnyDay, 2 RainDance, 3 Sandstorm
    30     1     n_act             number of legal actions (prefix mask)
    31     1     side_cond         carried, NOT encoded
    32     108   act[9] x 12 B     {u8 flags(bit0 switch, bit1 force_switch), u8 species,
                                    u8 item, u8 n_moves, u16 move[4]}
    140    4     user_tag  u32     opaque to the kernels (battle id for the driver)

``pack_observation`` performs the same string lookups as
ChoiceToVec.transform (choice_to_vec.py:40-61) and
StateFeatureExtractor._transform_* (state_feature_extractor.py:113-197) and
raises the same exceptions on unknown names (KeyError) / missing active
(AssertionError) / zero denominators (ZeroDivisionError).
"""
from __future__ import annotations

from typing import Iterable, List, Optional, Sequence

import numpy as np

from .observation import (ActivePokeStatus, BattleStatus, PossibleAction, RLPolicyObservation, RANK_KEYS)
from .tables import (MAX_ACTIONS, MAX_MOVES, N_STATE, STATUS_CODE, STATUS_FNT, WEATHER_CODE, Tables, load_tables,
                     NV_CONDITIONS, WEATHERS)

PK_ACT_DTYPE = np.dtype([("flags", "u1"), ("species", "u1"), ("item", "u1"), ("n_moves", "u1"),
                         ("move", "<u2", (MAX_MOVES,))])
PK_OBS_DTYPE = np.dtype([
    ("remaining", "u1", (2,)), ("total", "u1", (2,)),
    ("hp_cur", "<u2", (2,)), ("hp_max", "<u2", (2,)),
    ("status", "u1", (2,)), ("boost", "i1", (2, 7)),
    ("opp_species", "u1"), ("weather", "u1"), ("n_act", "u1"), ("side_cond", "u1"),
    ("act", PK_ACT_DTYPE, (MAX_ACTIONS,)),
    ("user_tag", "<u4"),
])
assert PK_ACT_DTYPE.itemsize == 12 and PK_OBS_DTYPE.itemsize == 144
PK_OBS_BYTES = 144

FLAG_SWITCH = 1
FLAG_FORCE = 2


def _status_code(s: str) -> int:
    if s == "" or s is None:
        return 0
    return STATUS_CODE.get(s, STATUS_FNT)


def pack_state_into(rec, bs, tables: Tables) -> None:
    """The battle-state half of a record (offsets 0..30) from a BattleStatus-like object."""
    sides = (bs.side_friend, bs.side_opponent)
    for i, side in enumerate(sides):
        ss = bs.side_statuses[side]
        active = ss.active
        assert active is not None  # state_feature_extractor.py:133,148,161,177
        if ss.total_pokes == 0 or active.hp_max == 0:
            raise ZeroDivisionError("division by zero")  # the reference divides in Python (:122,:150)
        rec["remaining"][i] = ss.remaining_pokes
        rec["total"][i] = ss.total_pokes
        rec["hp_cur"][i] = active.hp_current
        rec["hp_max"][i] = active.hp_max
        rec["status"][i] = _status_code(active.status)
        for k, key in enumerate(RANK_KEYS):
            rec["boost"][i][k] = active.ranks[key]
    opp = bs.side_statuses[bs.side_opponent].active
    rec["opp_species"] = tables.species_name_to_idx[opp.species]  # KeyError like dex.get_pokedex_by_name
    rec["weather"] = WEATHER_CODE.get(bs.weather, 0)


def pack_into(rec, obs, tables: Tables, max_actions: int = MAX_ACTIONS) -> None:
    """Fill one record (a numpy void of PK_OBS_DTYPE) from an observation object."""
    pack_state_into(rec, obs.battle_status, tables)
    acts = obs.possible_actions
    if len(acts) > max_actions:
        # feature_extractor.py:49 would raise a numpy broadcast ValueError here
        raise ValueError(f"{len(acts)} possible actions exceed output_dim {max_actions}")
    rec["n_act"] = len(acts)
    for a_idx, pa in enumerate(acts):
        ar = rec["act"][a_idx]
        ar["flags"] = (FLAG_SWITCH if pa.switch else 0) | (FLAG_FORCE if pa.force_switch else 0)
        ar["species"] = tables.species_id_to_idx[pa.poke]
        moves = list(pa.allMoves) if pa.switch else [pa.move]
        if len(moves) > MAX_MOVES:
            raise ValueError("more than 4 moves on one action")
        ar["n_moves"] = len(moves)
        for m_idx, mv in enumerate(moves):
            ar["move"][m_idx] = tables.move_to_idx[mv]
        ar["item"] = tables.item_to_idx[pa.item]


def pack_observation(obs, tables: Optional[Tables] = None, max_actions: int = MAX_ACTIONS) -> np.ndarray:
    """One observation -> array of shape (1,) PK_OBS_DTYPE."""
    return pack_observations([obs], tables, max_actions)


def pack_observations(obs_list: Sequence, tables: Optional[Tables] = None,
                      max_actions: int = MAX_ACTIONS) -> np.ndarray:
    tables = tables or load_tables()
    out = np.zeros(len(obs_list), dtype=PK_OBS_DTYPE)
    for i, obs in enumerate(obs_list):
        pack_into(out[i], obs, tables, max_actions)
    return out


def unpack_observation(rec, tables: Optional[Tables] = None) -> RLPolicyObservation:
    """Inverse of the packer: rebuild observation objects from one record.

    Used to drive object-level code (the oracle port / the reference itself /
    the CPU baseline) from synthetic packed batches.
    """
    tables = tables or load_tables()
    bs = BattleStatus("p1", None)
    for i, side in enumerate(("p1", "p2")):
        ss = bs.side_statuses[side]
        ss.remaining_pokes = int(rec["remaining"][i])
        ss.total_pokes = int(rec["total"][i])
        code = int(rec["status"][i])
        status = "" if code == 0 else (NV_CONDITIONS[code - 1] if code <= 6 else "fnt")
        if i == 1:
            name = tables.species_names[int(rec["opp_species"])]
        else:
            # the friend's species is not part of the state block; use the first action's
            # species when it is a move action, otherwise any valid name
            name = tables.species_names[int(rec["act"][0]["species"])] if int(rec["n_act"]) else "Snorlax"
        ap = ActivePokeStatus(f"{side}a: {name}", name, 55, "M", int(rec["hp_cur"][i]), int(rec["hp_max"][i]), status)
        ap.ranks = {k: int(rec["boost"][i][j]) for j, k in enumerate(RANK_KEYS)}
        ss.active = ap
    w = int(rec["weather"])
    bs.weather = BattleStatus.WEATHER_NONE if w == 0 else WEATHERS[w - 1]
    acts: List[PossibleAction] = []
    n_move = 0
    for a_idx in range(int(rec["n_act"])):
        ar = rec["act"][a_idx]
        sw = bool(ar["flags"] & FLAG_SWITCH)
        moves = [tables.moves[int(m)] for m in ar["move"][: int(ar["n_moves"])]]
        if sw:
            key = f"switch {a_idx + 2}"
        else:
            n_move += 1
            key = f"move {n_move}"
        acts.append(PossibleAction(key, tables.species_ids[int(ar["species"])], None if sw else moves[0], sw,
                                   bool(ar["flags"] & FLAG_FORCE), moves, tables.items[int(ar["item"])]))
    return RLPolicyObservation(bs, {}, acts)


# --------------------------------------------------------------------------------------
# dense (C, A) features -> packed record (reference `resume=True` replay buffers)
# --------------------------------------------------------------------------------------
def _ratio_pair(r: np.float32, max_den: int):
    """Smallest (num, den), 1 <= den <= max_den, 0 <= num <= den, with float32(num / den) == r: the encoder only ever
    sees the ratio (state_feature_extractor.py:122,150), so any such pair encodes to the same bits."""
    den = np.arange(1, max_den + 1, dtype=np.int64)
    num = np.rint(np.float64(r) * den).astype(np.int64)
    ok = ((num / den).astype(np.float32) == np.float32(r)) & (num >= 0) & (num <= den)
    hit = np.flatnonzero(ok)
    if hit.size == 0:
        raise ValueError(f"{float(r)!r} is not a ratio of two integers with a denominator <= {max_den}")
    return int(num[hit[0]]), int(den[hit[0]])


def record_from_dense(feat: np.ndarray, mask: np.ndarray, tables: Optional[Tables] = None) -> np.ndarray:
    """Invert FeatureExtractor.transform (feature_extractor.py:40-52): a dense (C, A) float32 feature array and its
    (A,) mask -> a packed record that ENCODES TO THE SAME BITS.  What a reference checkpoint written with
    `save_state(resume=True)` keeps in its replay buffer (replay_buffer.py:7-15) is exactly such arrays.  One-hot blocks
    invert exactly; ratios are recovered as the smallest integer pair with the same float32 quotient; the opponent's
    species is any species with the same type set (only its types are encoded, state_feature_extractor.py:125-138).
    Raises ValueError for arrays that are not the image of an observation."""
    tables = tables or load_tables()
    feat = np.asarray(feat, dtype=np.float32)
    mask = np.asarray(mask)
    C, A = feat.shape
    if C != tables.n_channels or A > MAX_ACTIONS or mask.shape != (A,):
        raise ValueError(f"expected ({tables.n_channels}, A <= {MAX_ACTIONS}) features and an (A,) mask")
    rec = np.zeros(1, dtype=PK_OBS_DTYPE)[0]
    sv = feat[:N_STATE, 0]
    for i in range(2):
        rec["remaining"][i], rec["total"][i] = _ratio_pair(sv[i], 255)
        rec["hp_cur"][i], rec["hp_max"][i] = _ratio_pair(sv[19 + i], 65535)
        st = np.flatnonzero(sv[21 + 6 * i:27 + 6 * i])
        if st.size > 1:
            raise ValueError("more than one non-volatile status set")
        rec["status"][i] = 0 if st.size == 0 else int(st[0]) + 1
        for k in range(7):
            v = sv[33 + 7 * i + k]
            rank = int(round(float(v) * 12.0 - 6.0))
            if not -6 <= rank <= 6 or np.float32((rank + 6) / 12) != v:
                raise ValueError(f"{float(v)!r} is not a rank feature")
            rec["boost"][i][k] = rank
    tmask = np.uint32(sum(1 << t for t in np.flatnonzero(sv[2:19])))
    cand = np.flatnonzero(tables.species_type_mask == tmask)
    if cand.size == 0:
        raise ValueError("no species has the opponent's type set")
    rec["opp_species"] = int(cand[0])
    w = np.flatnonzero(sv[47:50])
    if w.size > 1:
        raise ValueError("more than one weather set")
    rec["weather"] = 0 if w.size == 0 else int(w[0]) + 1
    n_act = int(mask.sum())
    if not (np.all(mask[:n_act] == 1) and np.all(mask[n_act:] == 0)):
        raise ValueError("the action mask is not a prefix mask")
    rec["n_act"] = n_act
    ch = feat[N_STATE:]
    for c in range(n_act):
        col = ch[:, c]
        sp = np.flatnonzero(col[tables.CH_SPECIES:tables.CH_MOVE])
        mv = np.flatnonzero(col[tables.CH_MOVE:tables.CH_ITEM])
        it = np.flatnonzero(col[tables.CH_ITEM:])
        if sp.size != 1 or it.size != 1 or not 1 <= mv.size <= MAX_MOVES:
            raise ValueError(f"action column {c} is not the image of a PossibleAction")
        ar = rec["act"][c]
        ar["flags"] = (FLAG_SWITCH if col[tables.CH_SWITCH] != 0 else 0) | (FLAG_FORCE if col[tables.CH_FORCE] != 0 else 0)
        ar["species"], ar["item"], ar["n_moves"] = int(sp[0]), int(it[0]), int(mv.size)
        ar["move"][:mv.size] = mv
    return rec


# --------------------------------------------------------------------------------------
# synthetic batches (SURVEY.md §8 d "Synthetic value distribution")
# --------------------------------------------------------------------------------------

def _distinct_moves(rng: np.random.Generator, shape, n_moves: int) -> np.ndarray:
    """``shape + (4,)`` move ids, distinct within each group of 4."""
    keys = rng.random(shape + (n_moves,), dtype=np.float32) if int(np.prod(shape)) * n_moves <= (1 << 24) else None
    if keys is not None:
        return np.argpartition(keys, MAX_MOVES, axis=-1)[..., :MAX_MOVES].astype(np.uint16)
    # large batches: rejection-free offset trick (sorted distinct sample of 4 out of n)
    base = np.sort(rng.integers(0, n_moves - 3, size=shape + (MAX_MOVES,)), axis=-1)
    out = (base + np.arange(MAX_MOVES)).astype(np.uint16)
    return rng.permuted(out, axis=-1)


def synth_packed(n: int, seed: int = 0, party_size: int = 6, kind: str = "full",
                 tables: Optional[Tables] = None) -> np.ndarray:
    """Seeded synthetic packed observations.

    kind="full":   4 move actions + (party_size-1) switch actions, every slot legal
                   (the benchmark distribution).
    kind="ragged": n_act ~ U{2..A}, ~15 % forced-switch states (switch actions only,
                   force flag set), fainted actives (0/100, status fnt), 0-3 move lists
                   shorter than 4 -- exercises masks and every status/weather code.
    """
    tables = tables or load_tables()
    rng = np.random.default_rng(seed)
    A = 4 + party_size - 1
    rec = np.zeros(n, dtype=PK_OBS_DTYPE)
    rec["total"][:] = party_size
    rec["remaining"] = rng.integers(1, party_size + 1, size=(n, 2))
    rec["hp_max"] = rng.integers(100, 401, size=(n, 2))
    rec["hp_cur"] = (rng.random((n, 2)) * rec["hp_max"]).astype(np.uint16) + 1
    rec["hp_cur"] = np.minimum(rec["hp_cur"], rec["hp_max"])
    st = rng.integers(0, 9, size=(n, 2))  # 0,1,2 -> none ; 3..8 -> cond 1..6
    rec["status"] = np.where(st < 3, 0, st - 2)
    rec["boost"] = rng.integers(-6, 7, size=(n, 2, 7))
    rec["opp_species"] = rng.integers(0, tables.n_species, size=n)
    rec["weather"] = rng.integers(0, 4, size=n)
    rec["user_tag"] = np.arange(n, dtype=np.uint32)

    act = rec["act"]
    active_species = rng.integers(0, tables.n_species, size=n)
    active_item = rng.integers(0, tables.n_items, size=n)
    active_moves = _distinct_moves(rng, (n,), tables.n_moves)            # (n,4)
    bench_species = rng.integers(0, tables.n_species, size=(n, party_size - 1))
    bench_item = rng.integers(0, tables.n_items, size=(n, party_size - 1))
    bench_moves = _distinct_moves(rng, (n, party_size - 1), tables.n_moves)  # (n,P-1,4)

    if kind == "full":
        for k in range(4):
            act["flags"][:, k] = 0
            act["species"][:, k] = active_species
            act["item"][:, k] = active_item
            act["n_moves"][:, k] = 1
            act["move"][:, k, 0] = active_moves[:, k]
        for j in range(party_size - 1):
            k = 4 + j
            act["flags"][:, k] = FLAG_SWITCH
            act["species"][:, k] = bench_species[:, j]
            act["item"][:, k] = bench_item[:, j]
            act["n_moves"][:, k] = 4
            act["move"][:, k, :] = bench_moves[:, j, :]
        rec["n_act"] = A
        return rec

    assert kind == "ragged"
    forced = rng.random(n) < 0.15
    if party_size == 1:
        forced[:] = False  # nothing to switch to
    fainted = rng.random((n, 2)) < 0.1
    fainted[:, 0] |= forced  # forced switch happens when the friend's active fainted
    rec["hp_cur"] = np.where(fainted, 0, rec["hp_cur"])
    rec["hp_max"] = np.where(fainted, 100, rec["hp_max"])  # battle_status.py:18-21
    rec["status"] = np.where(fainted, STATUS_FNT, rec["status"])
    n_moves_avail = rng.integers(1, 5, size=n)           # enabled moves of the active
    n_bench_avail = rng.integers(0, party_size, size=n)  # non-fainted bench
    short = rng.integers(1, 5, size=(n, party_size - 1))  # bench move-list lengths
    short = np.where(rng.random((n, party_size - 1)) < 0.8, 4, short)
    for i in range(n):
        k = 0
        if not forced[i]:
            for m in range(int(n_moves_avail[i])):
                a = act[i, k]
                a["flags"] = 0
                a["species"] = active_species[i]
                a["item"] = active_item[i]
                a["n_moves"] = 1
                a["move"][0] = active_moves[i, m]
                k += 1
        nb = int(n_bench_avail[i])
        if forced[i]:
            nb = max(nb, 1)
        nb = min(nb, party_size - 1)
        for j in range(nb):
            a = act[i, k]
            a["flags"] = FLAG_SWITCH | (FLAG_FORCE if forced[i] else 0)
            a["species"] = bench_species[i, j]
            a["item"] = bench_item[i, j]
            a["n_moves"] = short[i, j]
            a["move"][: short[i, j]] = bench_moves[i, j, : short[i, j]]
            k += 1
        rec["n_act"][i] = k
    return rec
