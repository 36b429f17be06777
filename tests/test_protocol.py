"""Request -> packed path against the reference's own get_possible_actions
(tools/make_golden_protocol.py ran it in the build container; the GPU box has no reference)."""
import copy
import gzip
import json
from collections import namedtuple

import numpy as np
import pytest

from conftest import GOLDEN


@pytest.fixture(scope="module")
def cases():
    with gzip.open(GOLDEN / "protocol_requests.json.gz", "rt") as f:
        return json.load(f)


@pytest.fixture(scope="module")
def statuses():
    """real battle states from the recorded battle log, to sit beside the synthetic requests"""
    from pokeai_b200.observation import battle_status_from_json
    with gzip.open(GOLDEN / "fixture_decisions.json.gz", "rt") as f:
        d = json.load(f)
    return [battle_status_from_json(x["battle_status"]) for x in d]


def test_pack_request_equals_possible_actions_route(cases, statuses, tables):
    """bytes of pack_request == bytes of pack_observation(RLPolicyObservation(bs, request, reference's list)),
    and the keys are the reference's simulator keys in order"""
    from pokeai_b200.observation import RLPolicyObservation, possible_actions_from_json
    from pokeai_b200.packed import pack_observation
    from pokeai_b200.protocol import pack_request, pack_requests
    n_checked = 0
    for i, c in enumerate(cases):
        bs = statuses[i % len(statuses)]
        pas = possible_actions_from_json(c["possible_actions"])
        rec, keys = pack_request(bs, c["request"], tables)
        assert keys == [pa.simulator_key for pa in pas], i
        want = pack_observation(RLPolicyObservation(bs, c["request"], pas), tables)
        assert rec.tobytes() == want.tobytes(), i
        n_checked += 1
    assert n_checked == len(cases) >= 200
    # batched form
    pairs = [(statuses[i % len(statuses)], c["request"]) for i, c in enumerate(cases)]
    recs, keys = pack_requests(pairs, tables)
    assert recs.shape == (len(cases),) and all(int(r["n_act"]) == len(k) for r, k in zip(recs, keys))
    assert recs[7].tobytes() == pack_request(*pairs[7], tables)[0].tobytes()


def test_struct_packer_writes_the_same_bytes(cases, statuses, tables):
    """pack_request_into (struct.pack_into, the lock-step driver's packer) == pack_request, byte for byte"""
    from pokeai_b200.packed import PK_OBS_BYTES
    from pokeai_b200.protocol import pack_request, pack_request_into
    buf = bytearray(b"\xff" * (3 * PK_OBS_BYTES))
    for i, c in enumerate(cases):
        bs = statuses[i % len(statuses)]
        rec, keys = pack_request(bs, c["request"], tables)
        got = pack_request_into(buf, PK_OBS_BYTES, bs, c["request"], tables)
        assert got == keys, i
        assert bytes(buf[PK_OBS_BYTES:2 * PK_OBS_BYTES]) == rec.tobytes(), i
        assert buf[:PK_OBS_BYTES] == b"\xff" * PK_OBS_BYTES and buf[2 * PK_OBS_BYTES:] == b"\xff" * PK_OBS_BYTES
    with pytest.raises(ValueError):
        pack_request_into(buf, 0, statuses[0], [c for c in cases if len(c["possible_actions"]) > 3][0]["request"], tables, 3)
    with pytest.raises(KeyError):
        bad = copy.deepcopy(cases[0]["request"])
        bad["side"]["pokemon"][0]["details"] = "Missingno, L50"
        pack_request_into(buf, 0, statuses[0], bad, tables)


def test_c_packer_writes_the_same_bytes_and_raises_the_same(cases, statuses, tables):
    """csrc/pkpack.c (the lock-step driver's packer when built) == pack_request_into_py: bytes, keys, exception types"""
    from pokeai_b200 import build as b
    from pokeai_b200 import protocol as P
    from pokeai_b200.packed import PK_OBS_BYTES
    if b.build_packer() is None:
        pytest.skip("Python headers or gcc not available")
    cpack = P._load_c_packer()
    assert cpack is not None

    def c_call(buf, off, bs, req, max_actions=9):
        return cpack(buf, off, bs, req, tables.species_name_to_idx, tables.move_to_idx, tables.item_to_idx, max_actions,
                     P.STATUS_CODE, P.STATUS_FNT, P.WEATHER_CODE)
    for i, c in enumerate(cases):
        bs = statuses[i % len(statuses)]
        b1, b2 = bytearray(b"\xaa" * (3 * PK_OBS_BYTES)), bytearray(b"\xaa" * (3 * PK_OBS_BYTES))
        k1 = P.pack_request_into_py(b1, PK_OBS_BYTES, bs, c["request"], tables)
        k2 = c_call(b2, PK_OBS_BYTES, bs, c["request"])
        assert k1 == k2 and b1 == b2, i
    buf = bytearray(PK_OBS_BYTES)
    big = [c for c in cases if len(c["possible_actions"]) > 3][0]["request"]
    bad = copy.deepcopy(cases[0]["request"])
    bad["side"]["pokemon"][0]["details"] = "Missingno, L50"
    no_active = {k: v for k, v in cases[0]["request"].items() if k != "active"}
    zero = copy.deepcopy(statuses[0])
    zero.side_statuses[zero.side_friend].active.hp_max = 0
    for exc, bs, req, kw in ((ValueError, statuses[0], big, {"max_actions": 3}), (KeyError, statuses[0], bad, {}),
                             (KeyError, statuses[0], no_active, {}), (ZeroDivisionError, zero, cases[0]["request"], {})):
        if req is no_active and req.get("forceSwitch"):
            continue
        with pytest.raises(exc):
            P.pack_request_into_py(buf, 0, bs, req, tables, **kw)
        with pytest.raises(exc):
            c_call(buf, 0, bs, req, **kw)
    with pytest.raises(ValueError):
        c_call(bytearray(100), 0, statuses[0], cases[0]["request"])      # the record does not fit the buffer
    assert P.pack_request_into is not None


def test_get_possible_actions_equals_reference(cases, tables):
    """the restated get_possible_actions against the reference's outputs, field by field"""
    from pokeai_b200.protocol import get_possible_actions
    for i, c in enumerate(cases):
        got = [pa._asdict() for pa in get_possible_actions(None, c["request"], tables)]
        assert got == c["possible_actions"], i


def test_pack_request_errors(cases, statuses, tables):
    from pokeai_b200.protocol import pack_request
    bs = statuses[0]
    turn = next(c["request"] for c in cases if c["request"].get("active") and len(c["request"]["side"]["pokemon"]) == 6)
    r = copy.deepcopy(turn)
    del r["active"]
    with pytest.raises(KeyError):  # common.py:135 request['active']
        pack_request(bs, r, tables)
    r = copy.deepcopy(turn)
    r["side"]["pokemon"][1]["details"] = "Missingno, L50"
    with pytest.raises(KeyError):  # dex.get_pokedex_by_name
        pack_request(bs, r, tables)
    r = copy.deepcopy(turn)
    r["side"]["pokemon"][0]["item"] = "notanitem"
    with pytest.raises(KeyError):  # choice_to_vec.py item lookup
        pack_request(bs, r, tables)
    r = copy.deepcopy(turn)
    for m in r["side"]["pokemon"]:
        m["condition"] = "10/10"
    r["active"][0].pop("trapped", None)
    for m in r["active"][0]["moves"]:
        m["disabled"] = False
    n = len(r["active"][0]["moves"]) + 5
    with pytest.raises(ValueError):  # more actions than the extractor's output_dim
        pack_request(bs, r, tables, max_actions=n - 1)
    assert len(pack_request(bs, r, tables, max_actions=n)[1]) == n


Cfg = namedtuple("Cfg", "hp_ratio alive_ratio only_opponent offset_at_end")


def _side(real, hp, alive):
    """a real side status with the two ratio methods of the reference's SideStatus (battle_status.py:132-150)"""
    ss = copy.copy(real)
    ss.get_mean_hp_ratio = lambda: hp
    ss.get_alive_ratio = lambda: alive
    return ss


class _FakeAgent:
    """records what the policy hands to act / stop_episode"""

    def __init__(self, output_dim):
        fe = type("FE", (), {})()
        fe.output_dim = output_dim
        self._feature_extractor = fe
        self.calls, self.stopped = [], []

    def act(self, rec, reward):
        self.calls.append((rec.copy(), reward))
        return int(rec["n_act"][0]) - 1

    def stop_episode(self, reward):
        self.stopped.append(reward)


def test_policy_rewards_and_single_choice(cases, statuses, tables):
    """surrogate reward = potential difference, first decision 0.0, single choices bypass the model and leave the
    potential untouched, offset_at_end subtracts the last potential (rl_policy.py:60-125)"""
    from pokeai_b200.protocol import PackedRLPolicy
    bs = copy.copy(statuses[0])
    real = bs.side_statuses
    bs.side_statuses = {bs.side_friend: _side(real[bs.side_friend], 0.8, 1.0),
                        bs.side_opponent: _side(real[bs.side_opponent], 0.5, 2 / 3)}
    multi = next(c["request"] for c in cases if len(c["possible_actions"]) >= 3)
    single = next(c["request"] for c in cases if len(c["possible_actions"]) == 1)
    agent = _FakeAgent(9)
    pol = PackedRLPolicy(agent, Cfg(1.0, 0.5, False, True), tables)
    pol.game_start()
    k0 = pol.choice_turn_start(bs, multi)
    p0 = (0.8 + 0.5) - (0.5 + 1 / 3)
    assert agent.calls[-1][1] == 0.0 and pol.last_reward_potential == pytest.approx(p0)
    assert k0 == [pa["simulator_key"] for pa in next(c for c in cases if c["request"] is multi)["possible_actions"]][-1]
    bs.side_statuses[bs.side_opponent] = _side(real[bs.side_opponent], 0.25, 1 / 3)
    assert pol.choice_force_switch(bs, single) in ("move 1", "move 2", "move 3", "move 4", "switch 1", "switch 2",
                                                   "switch 3", "switch 4", "switch 5", "switch 6")
    assert len(agent.calls) == 1 and pol.last_reward_potential == pytest.approx(p0)  # not consulted, not updated
    pol.choice_turn_start(bs, multi)
    p1 = (0.8 + 0.5) - (0.25 + 1 / 6)
    assert agent.calls[-1][1] == pytest.approx(p1 - p0)
    pol.game_end(1.0)
    assert agent.stopped == [pytest.approx(1.0 - p1)]
    pol2 = PackedRLPolicy(agent, Cfg(1.0, 0.0, True, False), tables)
    pol2.game_start()
    pol2.choice_turn_start(bs, multi)
    assert pol2.last_reward_potential == pytest.approx(-0.25)


def test_reward_potential_with_every_term_off_is_the_general_formula(statuses, tables):
    """the short cut for hp_ratio == alive_ratio == 0 returns the bits (sign of zero included) of rl_policy.py:60-74's
    formula, and keeps its state checks"""
    import struct
    from pokeai_b200.protocol import PackedRLPolicy
    agent = _FakeAgent(9)
    for only_opponent in (False, True):
        c = Cfg(0.0, 0.0, only_opponent, False)
        pol = PackedRLPolicy(agent, c, tables)
        for bs in statuses[:50]:
            sps = [ss.get_mean_hp_ratio() * c.hp_ratio + ss.get_alive_ratio() * c.alive_ratio
                   for ss in (bs.side_statuses[bs.side_friend], bs.side_statuses[bs.side_opponent])]
            want = -sps[1] if only_opponent else sps[0] - sps[1]
            assert struct.pack("d", pol._calc_reward_potential(bs)) == struct.pack("d", want)
    bad = copy.copy(statuses[0])
    bad.side_statuses = dict(bad.side_statuses)
    broken = copy.copy(bad.side_statuses[bad.side_opponent])
    broken.total_pokes = 0
    bad.side_statuses[bad.side_opponent] = broken
    with pytest.raises(AssertionError):
        pol._calc_reward_potential(bad)


@pytest.mark.gpu
def test_policy_on_gpu_matches_object_route(cases, statuses, tables):
    """PackedRLPolicy (single calls and choice_batch) picks what AgentVal.act picks on the reference-style
    observation objects"""
    import torch
    from pokeai_b200 import AgentVal, FeatureExtractor, MLPModel
    from pokeai_b200.observation import RLPolicyObservation, possible_actions_from_json
    from pokeai_b200.protocol import PackedRLPolicy
    dev = torch.device("cuda:0")
    sel = [c for c in cases if c["party_size"] == 6][:64]
    fe = FeatureExtractor(6)
    torch.manual_seed(5)
    model = MLPModel(fe.input_shape, fe.output_dim, n_layers=2, n_channels=64).to(dev)
    agent = AgentVal(model, fe)
    cfg = Cfg(0.0, 0.0, False, False)
    want, pols, bss, reqs = [], [], [], []
    for i, c in enumerate(sel):
        bs = statuses[(3 * i) % len(statuses)]
        pas = possible_actions_from_json(c["possible_actions"])
        if len(pas) == 1:
            want.append(pas[0].simulator_key)
        else:
            want.append(pas[agent.act(RLPolicyObservation(bs, c["request"], pas), 0.0)].simulator_key)
        pols.append(PackedRLPolicy(agent, cfg, tables))
        bss.append(bs)
        reqs.append(c["request"])
    single = [p.choice_turn_start(b, r) for p, b, r in zip(pols, bss, reqs)]
    assert single == want
    assert PackedRLPolicy.choice_batch(pols, bss, reqs) == want


@pytest.mark.parametrize("party_size", [1, 2, 3, 4, 5, 6])
def test_packers_agree_on_seeded_synthetic_requests(tables, party_size):
    """Wider than the 240 golden cases: 400 seeded (battle_status, request) pairs per party size from the lock-step
    driver's generator (forced switches, trapped, disabled moves, fainted bench, every rank / weather / condition) --
    pack_request (numpy fields), pack_request_into_py (struct) and csrc/pkpack.c give the same bytes and keys, and the
    record decodes through the packed oracle to the features of the object route."""
    import random
    from oracle import packed_oracle as PO
    from pokeai_b200 import build as b
    from pokeai_b200 import protocol as P
    from pokeai_b200.lockstep import synthetic_battle_status, synthetic_request
    from pokeai_b200.packed import PK_OBS_BYTES
    cpack = P._load_c_packer() if b.build_packer() is not None else None
    rng = random.Random(1000 + party_size)
    recs = []
    for i in range(400):
        bs = synthetic_battle_status(rng, tables, party_size)
        req = synthetic_request(rng, tables, party_size)
        rec, keys = P.pack_request(bs, req, tables)
        b1 = bytearray(b"\x55" * PK_OBS_BYTES)
        assert P.pack_request_into_py(b1, 0, bs, req, tables) == keys, i
        assert bytes(b1) == rec.tobytes(), i
        if cpack is not None:
            b2 = bytearray(b"\x55" * PK_OBS_BYTES)
            k2 = cpack(b2, 0, bs, req, tables.species_name_to_idx, tables.move_to_idx, tables.item_to_idx, 9,
                       P.STATUS_CODE, P.STATUS_FNT, P.WEATHER_CODE)
            assert k2 == keys and b2 == b1, i
        assert 1 <= len(keys) <= 9 and int(np.atleast_1d(rec)["n_act"][0]) == len(keys)
        recs.append(rec)
    # the records are well formed for the decision path: the packed oracle encodes them, the mask is the action count
    packed = np.concatenate([np.atleast_1d(r) for r in recs])
    feats, mask = PO.encode(packed, 6, tables)
    assert feats.shape == (400, tables.n_channels, 9) and np.isfinite(feats).all()
    assert (mask.sum(axis=1) == packed["n_act"]).all()
