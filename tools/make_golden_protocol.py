"""Golden vectors for the request -> packed path, produced by the REAL reference
get_possible_actions (pokeai/ai/common.py:143-174) in the build container.

    PYTHONDONTWRITEBYTECODE=1 python tools/make_golden_protocol.py

tests/golden/protocol_requests.json.gz: seeded synthetic `|request|` objects in the simulator's format
(turn start, forced switch, trapped, disabled moves, fainted bench members, Hidden Power / Return / Frustration
renames, parties of 1..6) and, for each, the reference's PossibleAction list.
"""
import gzip
import json
import os
import random
import sys
from pathlib import Path

sys.dont_write_bytecode = True
ROOT = Path(__file__).resolve().parents[1]
REF = Path(os.environ.get("POKEAI_REFERENCE", "/root/reference"))
sys.path.insert(0, str(ROOT))
sys.path.insert(1, str(REF))

from pokeai.ai.common import get_possible_actions  # noqa: E402
from pokeai.ai.dex import dex  # noqa: E402
from pokeai.util import DATASET_DIR, json_load  # noqa: E402

poke_ids = json_load(DATASET_DIR.joinpath("all_pokemons.json"))
move_ids = [m for m in json_load(DATASET_DIR.joinpath("all_moves_with_hiddenpower_type.json")) if m != "hiddenpower"]
item_ids = json_load(DATASET_DIR.joinpath("all_items.json"))
id2name = {dex._pokedex[i]["id"]: n for n, i in dex._poke2id.items()}


def side_move(rng, mid):
    """a move id as request['side']['pokemon'][i]['moves'] spells it"""
    if mid == "return":
        return "return102"
    if mid == "frustration":
        return "frustration1"
    return mid


def active_move(rng, mid):
    """{move, id, disabled} as request['active'][0]['moves'][j] spells it"""
    if mid.startswith("hiddenpower") and mid != "hiddenpower":
        t = mid[len("hiddenpower"):]
        return {"move": f"Hidden Power {t.capitalize()} 70", "id": "hiddenpower", "pp": 24, "maxpp": 24,
                "target": "normal", "disabled": False}
    name = {"return": "Return 102", "frustration": "Frustration 1"}.get(mid, mid.capitalize())
    return {"move": name, "id": mid, "pp": 10, "maxpp": 16, "target": "normal", "disabled": False}


def make_request(rng, size):
    species = rng.sample(poke_ids, size)
    kind = rng.choice(["turn", "turn", "turn", "force", "trapped"])
    active_idx = 0
    members = []
    for i, sp in enumerate(species):
        moves = rng.sample(move_ids, rng.randint(1, 4))
        hp_max = rng.randint(100, 400)
        fainted = i != active_idx and rng.random() < 0.25
        if kind == "force" and i == active_idx:
            fainted = True
        cond = "0 fnt" if fainted else f"{rng.randint(1, hp_max)}/{hp_max}" + rng.choice(["", "", " par", " slp"])
        members.append({"ident": f"p1: {id2name[sp]}", "details": f"{id2name[sp]}, L{rng.randint(50, 55)}" +
                        rng.choice(["", ", M", ", F"]), "condition": cond, "active": i == active_idx,
                        "stats": {"atk": 100, "def": 100, "spa": 100, "spd": 100, "spe": 100},
                        "moves": [side_move(rng, m) for m in moves], "baseAbility": "noability",
                        "item": rng.choice(item_ids + ["", ""]), "pokeball": "pokeball", "_moves": moves})
    req = {"side": {"name": "p1", "id": "p1", "pokemon": members}, "rqid": rng.randint(1, 99)}
    if kind == "force":
        req["forceSwitch"] = [True]
    else:
        am = [active_move(rng, m) for m in members[active_idx]["_moves"]]
        for m in am:
            if len(am) > 1 and rng.random() < 0.15:
                m["disabled"] = True
        act = {"moves": am}
        if kind == "trapped":
            act["trapped"] = True
        req["active"] = [act]
    for m in members:
        del m["_moves"]
    return req


def main():
    rng = random.Random(20240611)
    cases = []
    for n in range(240):
        size = rng.choice([1, 2, 3, 3, 4, 6, 6, 6])
        req = make_request(rng, size)
        pas = get_possible_actions(None, req)  # the battle status argument is not read (common.py:143-174)
        cases.append({"party_size": size, "request": req, "possible_actions": [pa._asdict() for pa in pas]})
    out = ROOT / "tests" / "golden" / "protocol_requests.json.gz"
    with gzip.GzipFile(out, "wb", mtime=0) as f:
        f.write(json.dumps(cases, separators=(",", ":")).encode())
    kinds = {}
    for c in cases:
        k = "force" if c["request"].get("forceSwitch") else ("trapped" if c["request"]["active"][0].get("trapped") else "turn")
        kinds[k] = kinds.get(k, 0) + 1
    print(out, out.stat().st_size, "bytes", kinds, "max actions", max(len(c["possible_actions"]) for c in cases))


main()
