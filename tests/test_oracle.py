"""CPU: pin the oracle (oracle/ref_port.py, oracle/packed_oracle.py) and the host packer
against the golden vectors produced by the real reference (tools/make_golden.py)."""
import json

import numpy as np
import pytest
import torch

from conftest import GOLDEN, bits
from oracle import packed_oracle, ref_port
from pokeai_b200.observation import (RLPolicyObservation, battle_status_from_json, possible_actions_from_json)
from pokeai_b200.packed import PK_OBS_DTYPE, pack_observations, synth_packed, unpack_observation


def _fixture_obs(dec):
    return [RLPolicyObservation(battle_status_from_json(d["battle_status"]), {},
                                possible_actions_from_json(d["possible_actions"])) for d in dec]


def test_port_matches_reference_on_real_battle_log(fixture_golden, tables):
    dec, feat, mask = fixture_golden
    assert len(dec) == 220
    obs = _fixture_obs(dec)
    f, m = ref_port.transform_batch(obs, 3, tables)
    assert np.array_equal(bits(f), bits(feat)) and np.array_equal(m, mask) and m.dtype == np.int32


def test_logged_actions_are_masked_argmax(fixture_golden):
    """agent.py:34-35 -- the logged action is argmax of the logged (already masked) Q vector."""
    dec, _, mask = fixture_golden
    q = np.array([[np.float32(x) for x in d["q_func"]] for d in dec], dtype=np.float32)
    assert np.array_equal(np.isneginf(q), mask == 0)
    act = ref_port.masked_argmax(q, mask)
    assert np.array_equal(act, np.array([d["action"] for d in dec]))
    for d, a in zip(dec, act):
        assert d["possible_actions"][a]["simulator_key"] == d["choice"]


def test_packer_roundtrip_and_packed_oracle_on_real_battle_log(fixture_golden, tables):
    dec, feat, mask = fixture_golden
    rec = pack_observations(_fixture_obs(dec), tables, max_actions=6)
    f, m = packed_oracle.encode(rec, 3, tables)
    assert np.array_equal(bits(f), bits(feat)) and np.array_equal(m, mask)


@pytest.mark.parametrize("name", ["full6", "ragged6", "ragged3", "ragged1", "full4"])
def test_oracles_match_reference_on_synthetic(synth_golden, tables, name):
    rec, ps, feat, mask = synth_golden[name]
    obs = [unpack_observation(r, tables) for r in rec]
    f, m = ref_port.transform_batch(obs, ps, tables)
    assert np.array_equal(bits(f), bits(feat)) and np.array_equal(m, mask)
    f2, m2 = packed_oracle.encode(rec, ps, tables)
    assert np.array_equal(bits(f2), bits(feat)) and np.array_equal(m2, mask)
    # the packer inverts unpack (user_tag is not part of an observation)
    rec2 = pack_observations(obs, tables, 4 + ps - 1)
    ref = rec.copy()
    ref["user_tag"] = 0
    assert rec2.tobytes() == ref.tobytes()


def test_packed_oracle_equals_port_random(tables):
    for seed, ps, kind in [(101, 6, "ragged"), (102, 2, "ragged"), (103, 5, "full")]:
        rec = synth_packed(200, seed=seed, party_size=ps, kind=kind, tables=tables)
        f, m = packed_oracle.encode(rec, ps, tables)
        f2, m2 = ref_port.transform_batch([unpack_observation(r, tables) for r in rec], ps, tables)
        assert np.array_equal(bits(f), bits(f2)) and np.array_equal(m, m2)


def test_dim_meanings(tables):
    g = json.loads((GOLDEN / "dim_meanings.json").read_text())
    ours = ["state/" + m for m in tables.state_dim_meanings()] + ["choice/" + m for m in tables.choice_dim_meanings()]
    assert ours == g["party_size_6"]
    assert g["input_shape_6"] == [617, 9] and g["input_shape_3"] == [617, 6]
    assert tables.n_channels == 617


def _sd(z, prefix):
    return {k[len(prefix):]: torch.from_numpy(z[k]) for k in z.files if k.startswith(prefix)}


@pytest.mark.parametrize("name", ["default", "small"])
def test_mlp_port_matches_reference(synth_golden, name):
    z = np.load(GOLDEN / "mlp_forward.npz")
    rec, ps, feat, mask = synth_golden[str(z[f"{name}_src"])]
    n = int(z[f"{name}_n"])
    sd = _sd(z, f"{name}_sd_")
    with torch.no_grad():
        q = ref_port.mlp_forward(sd, torch.from_numpy(feat[:n])).numpy()
    np.testing.assert_allclose(q, z[f"{name}_q"], rtol=1e-6, atol=1e-6)


def test_dqn_update_port_matches_reference(tables):
    z = np.load(GOLDEN / "dqn_update.npz")
    ps = int(z["party_size"])
    rec_s = np.frombuffer(z["packed_s"].tobytes(), dtype=PK_OBS_DTYPE)
    rec_n = np.frombuffer(z["packed_n"].tobytes(), dtype=PK_OBS_DTYPE)
    fs, _ = packed_oracle.encode(rec_s, ps, tables)
    fn, mn = packed_oracle.encode(rec_n, ps, tables)
    sd = {k: v.clone() for k, v in _sd(z, "sd0_").items()}
    tgt = _sd(z, "tgt_")
    adam = {}
    for step in (1, 2):
        loss, grads = ref_port.dqn_update_grads(sd, tgt, fs, z["action"], z["reward"], fn, mn, ~z["terminal"],
                                                float(z["gamma"]))
        assert abs(loss - float(z[f"loss{step}"])) < 1e-6
        for k, g in grads.items():
            np.testing.assert_allclose(g.numpy(), z[f"grad{step}_{k}"], rtol=1e-5, atol=1e-7)
        ref_port.clamp_adam_step(sd, grads, adam, float(z["lr"]), step)
        for k, p in sd.items():
            np.testing.assert_allclose(p.numpy(), z[f"sd{step}_{k}"], rtol=1e-6, atol=1e-7)


@pytest.mark.parametrize("tag", ["a", "b"])
def test_bn_update_port_matches_reference(tables, tag):
    """The written-out train-mode BatchNorm update (ref_port.dqn_update_grads_bn) against the real reference's
    Trainer._update with bn: true (tools/make_golden_bn.py): loss, pre-clamp gradients, running statistics and counters
    of both networks, and the parameters after clamp + Adam wherever the step is well conditioned (|g| >= 1e-6: at the
    first steps Adam moves a weight by lr * g / (|g| + 1e-8), so a gradient that is zero up to rounding gives a step
    that is noise in any implementation, torch's own included)."""
    z = np.load(GOLDEN / "dqn_update_bn.npz")
    p = tag + "_"
    ps = int(z[p + "party_size"])
    rec_s = np.frombuffer(z[p + "packed_s"].tobytes(), dtype=PK_OBS_DTYPE)
    rec_n = np.frombuffer(z[p + "packed_n"].tobytes(), dtype=PK_OBS_DTYPE)
    fs, _ = packed_oracle.encode(rec_s, ps, tables)
    fn, mn = packed_oracle.encode(rec_n, ps, tables)
    sd = {k: v.clone() for k, v in _sd(z, p + "sd0_").items()}
    tgt = {k: v.clone() for k, v in _sd(z, p + "tgt0_").items()}
    is_param = lambda k: not k.endswith(("running_mean", "running_var", "num_batches_tracked"))
    adam = {}
    shaky = {}
    for step in (1, 2):
        loss, grads = ref_port.dqn_update_grads_bn(sd, tgt, fs, z[p + "action"], z[p + "reward"], fn, mn, ~z[p + "terminal"],
                                                   float(z[p + "gamma"]), double_dqn=bool(z[p + "double_dqn"]))
        assert abs(loss - float(z[f"{p}loss{step}"])) < 1e-6
        for k, g in grads.items():
            want = z[f"{p}grad{step}_{k}"]
            np.testing.assert_allclose(g.numpy(), want, rtol=1e-4, atol=2e-7, err_msg=k)
            shaky[k] = shaky.get(k, False) | ((np.abs(want) < 1e-6) & (want != 0))
        params = {k: v for k, v in sd.items() if is_param(k)}
        ref_port.clamp_adam_step(params, grads, adam, float(z[p + "lr"]), step)
        for k, v in sd.items():
            want = z[f"{p}sd{step}_{k}"]
            if k.endswith("num_batches_tracked"):
                assert int(v) == int(want), k
            elif is_param(k):
                ok = ~shaky[k]
                np.testing.assert_allclose(v.numpy()[ok], want[ok], rtol=1e-5, atol=2e-6, err_msg=k)
            else:
                np.testing.assert_allclose(v.numpy(), want, rtol=1e-5, atol=1e-6, err_msg=k)
        for k, v in tgt.items():
            if "bn_layers" in k and not is_param(k):
                want = z[f"{p}tgt{step}_{k}"]
                if k.endswith("num_batches_tracked"):
                    assert int(v) == int(want), k
                else:
                    np.testing.assert_allclose(v.numpy(), want, rtol=1e-5, atol=1e-6, err_msg="target " + k)


def test_select_restatements():
    rng = np.random.default_rng(0)
    q = rng.standard_normal((64, 9)).astype(np.float32)
    mask = (rng.random((64, 9)) < 0.6).astype(np.int32)
    mask[:, 0] |= (mask.sum(1) == 0)
    q[3, :] = 0.5  # ties -> first legal max
    a = ref_port.masked_argmax(q, mask)
    assert a[3] == np.flatnonzero(mask[3])[0]
    assert all(mask[b, a[b]] == 1 for b in range(64))
    u1, u2 = rng.random(64).astype(np.float32), rng.random(64).astype(np.float32)
    e = ref_port.epsilon_greedy(q, mask, 0.3, u1, u2)
    assert np.array_equal(e[u1 >= 0.3], a[u1 >= 0.3]) and all(mask[b, e[b]] == 1 for b in range(64))
    s, p = ref_port.softmax_sample(q, mask, 1.0, u2)
    assert all(mask[b, s[b]] == 1 for b in range(64))
    np.testing.assert_allclose(p.sum(1), 1.0, rtol=1e-5)
    assert (p[mask == 0] == 0).all()


@pytest.mark.parametrize("name", ["full6", "ragged6", "ragged3", "ragged1", "full4"])
def test_dense_features_invert_to_equivalent_records(synth_golden, tables, name):
    """reference `resume=True` replay buffers hold dense (617, A) arrays: record_from_dense must give records that
    encode back to the same bits (the reference's own FeatureExtractor outputs are the input here)"""
    from oracle import packed_oracle
    from pokeai_b200.packed import PK_OBS_DTYPE, record_from_dense
    rec, ps, feat, mask = synth_golden[name]
    out = np.zeros(len(feat), dtype=PK_OBS_DTYPE)
    for i in range(len(feat)):
        out[i] = record_from_dense(feat[i], mask[i], tables)
    f2, m2 = packed_oracle.encode(out, ps, tables)
    assert np.array_equal(bits(f2), bits(feat)) and np.array_equal(m2, mask)
    with pytest.raises(ValueError):
        bad = feat[0].copy()
        bad[19, :] = 1.5  # hp_current > hp_max: not the image of an observation
        record_from_dense(bad, mask[0], tables)


def test_replay_buffer_accepts_reference_items(synth_golden, tables):
    """a reference ReplayBuffer (deque of dense ReplayBufferItem) converts into the packed ring"""
    from collections import deque, namedtuple
    from oracle import packed_oracle
    from pokeai_b200.replay_buffer import from_reference
    rec, ps, feat, mask = synth_golden["ragged6"]
    Item = namedtuple("ReplayBufferItem", "state action_mask action next_state next_action_mask reward")

    class RefBuffer:  # replay_buffer.py:18-32
        def __init__(self):
            self.buffer = deque(maxlen=1000)

    rb = RefBuffer()
    for i in range(0, 40, 2):
        last = i % 8 == 6
        rb.buffer.append(Item(feat[i], mask[i], i % 4, None if last else feat[i + 1], None if last else mask[i + 1], 0.5 * i))
    ours = from_reference(rb)
    assert len(ours) == 20 and ours.maxlen == 1000
    s, nx, act, rew, term = ours.gather(np.arange(20))
    fs, ms = packed_oracle.encode(s, ps, tables)
    assert np.array_equal(bits(fs), bits(feat[0:40:2])) and np.array_equal(ms, mask[0:40:2])
    keep = term == 0
    fn, mn = packed_oracle.encode(nx[keep], ps, tables)
    assert np.array_equal(bits(fn), bits(feat[1:40:2][keep])) and int(term.sum()) == 5
    assert list(act) == [i % 4 for i in range(0, 40, 2)] and rew[3] == 3.0
