"""GPU parity of the drop-in host classes (FeatureExtractor / MLPModel / AgentVal / AgentTrain)
against the golden vectors of the real reference and the oracle.  Everything goes through the C ABI."""
import numpy as np
import pytest
import torch

from conftest import GOLDEN, bits

pytestmark = pytest.mark.gpu

Q_RTOL, Q_ATOL = 1e-5, 1e-5


def _fixture_obs(dec):
    from pokeai_b200.observation import RLPolicyObservation, battle_status_from_json, possible_actions_from_json
    return [RLPolicyObservation(battle_status_from_json(d["battle_status"]), {},
                                possible_actions_from_json(d["possible_actions"])) for d in dec]


def _model_from(sd, fe):
    from pokeai_b200.mlp_model import MLPModel
    nl = sum(1 for k in sd if k.startswith("layers.") and k.endswith(".weight"))
    H = sd["layers.0.weight"].shape[0]
    m = MLPModel(fe.input_shape, fe.output_dim, n_layers=nl, n_channels=H)
    m.load_state_dict(sd)
    return m.eval()


def test_feature_extractor_matches_reference_on_real_log(fixture_golden):
    from pokeai_b200.feature_extractor import FeatureExtractor
    dec, feat, mask = fixture_golden
    obs = _fixture_obs(dec)
    fe = FeatureExtractor(3)
    assert fe.input_shape == (617, 6) and fe.output_dim == 6
    import json
    assert fe.get_dim_meanings() == json.loads((GOLDEN / "dim_meanings.json").read_text())["party_size_6"]
    f1, m1 = fe.transform(obs[0])
    assert f1.dtype == np.float32 and f1.shape == (617, 6) and m1.dtype == np.int32 and m1.shape == (6,)
    assert np.array_equal(bits(f1), bits(feat[0])) and np.array_equal(m1, mask[0])
    fb, mb = fe.transform_batch(obs)  # grows the session past its initial capacity
    assert np.array_equal(bits(fb), bits(feat)) and np.array_equal(mb, mask)
    # the sub-extractors the reference exposes (party_evaluator_quick.py:48,53)
    st = fe.state_feature_extractor.transform(obs[5])
    ch = fe.choice_to_vec.transform(obs[5])
    assert np.array_equal(bits(st), bits(feat[5][:50, 0]))
    assert np.array_equal(bits(ch), bits(feat[5][50:, :len(obs[5].possible_actions)]))


def test_agent_val_on_real_log(fixture_golden, tables):
    """AgentVal.act on the 220 logged decisions == masked argmax of the oracle's Q for the same weights;
    act_batch == act one by one."""
    from oracle import ref_port
    from pokeai_b200.agent import AgentVal
    from pokeai_b200.feature_extractor import FeatureExtractor
    dec, feat, mask = fixture_golden
    obs = _fixture_obs(dec)
    fe = FeatureExtractor(3)
    sd = ref_port.init_mlp_params(617, 3, 64, seed=11)
    agent = AgentVal(_model_from(sd, fe), fe)
    with torch.no_grad():
        qo = ref_port.mlp_forward(sd, torch.from_numpy(feat)).numpy()
    want = ref_port.masked_argmax(qo, mask)
    got_batch = agent.act_batch(obs)
    q, _ = agent.q_and_action_batch(obs)
    np.testing.assert_allclose(q, qo, rtol=Q_RTOL, atol=Q_ATOL)
    assert np.array_equal(got_batch, ref_port.masked_argmax(q, mask))
    # where the oracle's top two legal Q differ by more than the tolerance the action must be the oracle's
    qm = np.where(mask == 1, qo, -np.inf)
    top2 = np.sort(qm, axis=1)[:, -2:]
    clear = (top2[:, 1] - top2[:, 0]) > 1e-4
    assert clear.sum() > 150 and np.array_equal(got_batch[clear], want[clear])
    for i in (0, 7, 100, 219):
        assert agent.act(obs[i], 0.0) == int(got_batch[i])
    agent.stop_episode(0.0)


def test_agent_dense_q_path(synth_golden, tables):
    """_calc_q_vector[_batch] on dense arrays (party_evaluator.py:102) == oracle forward."""
    from oracle import ref_port
    from pokeai_b200.agent import AgentVal
    from pokeai_b200.feature_extractor import FeatureExtractor
    rec, ps, feat, mask = synth_golden["ragged6"]
    fe = FeatureExtractor(ps)
    sd = ref_port.init_mlp_params(617, 2, 64, seed=5)
    agent = AgentVal(_model_from(sd, fe), fe)
    with torch.no_grad():
        qo = ref_port.mlp_forward(sd, torch.from_numpy(feat)).numpy()
    np.testing.assert_allclose(agent._calc_q_vector_batch(feat), qo, rtol=Q_RTOL, atol=Q_ATOL)
    np.testing.assert_allclose(agent._calc_q_vector(feat[3]), qo[3], rtol=Q_RTOL, atol=Q_ATOL)
    a = agent._act_by_model(feat[3], mask[3])
    assert a == int(ref_port.masked_argmax(agent._calc_q_vector_batch(feat[3:4]), mask[3:4])[0])


def test_agent_train_follows_the_reference_rng_stream(tables):
    """AgentTrain.act consumes numpy's global RNG exactly like agent_train.py:30-33, so a seeded episode
    explores at the same steps and picks the same random slots as the reference would."""
    from oracle import ref_port
    from pokeai_b200.agent import AgentTrain
    from pokeai_b200.feature_extractor import FeatureExtractor
    from pokeai_b200.packed import synth_packed, unpack_observation
    from oracle import packed_oracle
    rec = synth_packed(60, seed=9, party_size=6, kind="ragged", tables=tables)
    obs = [unpack_observation(r, tables) for r in rec]
    fe = FeatureExtractor(6)
    sd = ref_port.init_mlp_params(617, 2, 64, seed=2)
    agent = AgentTrain(_model_from(sd, fe), fe, epsilon=0.4)
    fo, mo = packed_oracle.encode(rec, 6, tables)
    with torch.no_grad():
        qo = ref_port.mlp_forward(sd, torch.from_numpy(fo)).numpy()
    greedy = ref_port.masked_argmax(qo, mo)
    np.random.seed(123)
    got = [agent.act(o, 0.1 * i) for i, o in enumerate(obs)]
    agent.stop_episode(1.0)
    np.random.seed(123)
    want = []
    for i in range(len(obs)):  # the reference's control flow
        if np.random.random() < 0.4:
            want.append(int(np.random.choice(np.flatnonzero(mo[i]))))
        else:
            want.append(int(greedy[i]))
    assert got == want
    buf = agent._replay_buffer
    assert len(buf) == len(obs)  # one item per step; the last one terminal (agent_train.py:26-29,39-41)
    items = list(buf.buffer)
    assert items[-1].next_state is None and items[-1].reward == 1.0
    assert items[0].action == want[0] and items[0].reward == pytest.approx(0.1)
    assert bytes(items[0].next_state.tobytes()) == bytes(items[1].state.tobytes())
    assert np.array_equal(items[3].action_mask, mo[3])


def test_decision_pipeline_matches_blocking_calls(tables):
    """DecisionPipeline (pk_session_submit_host / pk_session_wait, two or three batches in flight) returns, in
    submission order, exactly what blocking act_batch calls return; ragged batch sizes; weight change mid-stream"""
    from pokeai_b200 import AgentVal, DecisionPipeline, FeatureExtractor, MLPModel
    from pokeai_b200.packed import synth_packed
    fe = FeatureExtractor(6)
    torch.manual_seed(3)
    model = MLPModel(fe.input_shape, fe.output_dim, n_layers=2, n_channels=64).to("cuda:0")
    agent = AgentVal(model, fe)
    sizes = [512, 1, 300, 512, 17, 128, 511]
    batches = [synth_packed(n, seed=40 + i, party_size=6, kind="ragged", tables=tables) for i, n in enumerate(sizes)]
    want = [agent.act_batch(b) for b in batches]
    wq = [agent.q_and_action_batch(b)[0].copy() for b in batches]
    for depth in (1, 2, 3):
        pipe = DecisionPipeline(agent, 512, depth=depth)
        got = list(pipe.map(batches))
        assert len(got) == len(want) and all(np.array_equal(g, w) for g, w in zip(got, want)), depth
        # by hand, with Q
        pipe.submit(batches[2])
        if depth > 1:
            pipe.submit(batches[4])
        else:
            with pytest.raises(RuntimeError):
                pipe.submit(batches[4])
        q, a = pipe.result(with_q=True)
        assert np.array_equal(a, want[2]) and np.array_equal(q, wq[2])
        if depth > 1:
            assert np.array_equal(pipe.result(), want[4])
        with pytest.raises(RuntimeError):
            pipe.result()
        with pytest.raises(ValueError):
            pipe.submit(synth_packed(513, seed=1, party_size=6, kind="full", tables=tables))
        pipe.close()
    # records handed over in pinned memory: read in place, same results; pageable arrays passed as "pinned" still work
    from pokeai_b200 import pinned_records
    pipe = DecisionPipeline(agent, 512, depth=2)
    pins = []
    for bt in batches:
        p_ = pinned_records(len(bt))
        p_.view(np.uint8)[...] = bt.view(np.uint8)
        pins.append(p_)
    got = list(pipe.map(pins, pinned=True))
    assert all(np.array_equal(g, w) for g, w in zip(got, want))
    got = list(pipe.map(batches, pinned=True))
    assert all(np.array_equal(g, w) for g, w in zip(got, want))
    with pytest.raises(ValueError):
        pipe.submit(np.zeros(4, dtype=np.uint8), pinned=True)
    pipe.close()
    # new weights while a batch is in flight: the batch in flight keeps the old ones, the next one sees the new
    pipe = DecisionPipeline(agent, 512, depth=2)
    pipe.submit(batches[0])
    with torch.no_grad():
        model.output.weight.mul_(-1.0)
    pipe.submit(batches[0])
    a_old, a_new = pipe.result(), pipe.result()
    assert np.array_equal(a_old, want[0])
    assert np.array_equal(a_new, agent.act_batch(batches[0])) and not np.array_equal(a_new, a_old)
    pipe.close()
