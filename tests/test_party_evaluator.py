"""Party evaluator against golden vectors produced by the reference's own PartyEvaluator /
PartyEvaluatorQuick (tools/make_golden_party.py).  The CPU tests check the packing (through the oracle encoder,
bit-exact features and masks); the GPU tests check Q and best-Q through the fused kernel."""
import gzip
import json

import numpy as np
import pytest
import torch

from conftest import GOLDEN, bits


@pytest.fixture(scope="module")
def gold():
    with gzip.open(GOLDEN / "party_eval.json.gz", "rt") as f:
        g = json.load(f)
    w = np.load(GOLDEN / "party_eval_weights.npz")
    return g, w


class _PackOnlyAgent:
    """enough of an Agent for PartyEvaluatorQuick.pack (no device needed)"""

    def __init__(self, output_dim):
        class _FE:
            pass
        self._feature_extractor = _FE()
        self._feature_extractor.output_dim = output_dim


@pytest.mark.parametrize("size", [3, 6])
def test_packed_party_observation_encodes_to_the_reference_features(gold, tables, size):
    from oracle import packed_oracle
    from pokeai_b200.party_evaluator import PartyEvaluatorQuick
    g, _ = gold
    G = g[str(size)]
    C, A = G["shape"]
    pe = PartyEvaluatorQuick(_PackOnlyAgent(A), size)
    for case in G["cases"]:
        rec = pe.pack(case["party"], G["opponents"])
        f, m = packed_oracle.encode(rec, size, tables)
        for i, sp in enumerate(case["features"]):
            ref = np.zeros(C * A, dtype=np.float32)
            ref[np.asarray(sp["idx"], dtype=np.int64)] = np.asarray(sp["val"], dtype=np.float32)
            assert np.array_equal(bits(f[i].ravel()), bits(ref))
            assert m[i].tolist() == case["masks"][i]


def test_pack_rejects_bad_parties(tables):
    from pokeai_b200.party_evaluator import PartyEvaluatorQuick
    pe = PartyEvaluatorQuick(_PackOnlyAgent(6), 3)
    with pytest.raises(AssertionError):
        pe.pack([{"species": "snorlax", "moves": ["tackle"], "item": ""}], ["miltank"])  # wrong party size
    party = [{"species": "snorlax", "moves": ["tackle"], "item": ""}] * 3
    with pytest.raises(KeyError):
        pe.pack(party, ["notapokemon"])
    with pytest.raises(ValueError):
        PartyEvaluatorQuick(_PackOnlyAgent(9), 3)


@pytest.mark.gpu
@pytest.mark.parametrize("size", [3, 6])
def test_party_q_matches_reference(gold, tables, size):
    from pokeai_b200.agent import AgentVal
    from pokeai_b200.feature_extractor import FeatureExtractor
    from pokeai_b200.mlp_model import MLPModel
    from pokeai_b200.party_evaluator import PartyEvaluatorQuick
    g, w = gold
    G = g[str(size)]
    dev = torch.device("cuda:0")
    fe = FeatureExtractor(size, device=dev)
    model = MLPModel(fe.input_shape, fe.output_dim, n_layers=2, n_channels=64).eval()
    model.load_state_dict({k[len(f"{size}_"):]: torch.from_numpy(w[k]) for k in w.files if k.startswith(f"{size}_")})
    pe = PartyEvaluatorQuick(AgentVal(model, fe), size)
    parties = [c["party"] for c in G["cases"]]
    for case in G["cases"]:
        for i, opp in enumerate(G["opponents"]):
            q = pe.calc_q_func(case["party"], opp)
            ref = np.array([x if x is not None else -np.inf for x in case["q"][i]], dtype=np.float32)
            assert np.array_equal(np.isinf(q), np.isinf(ref))
            np.testing.assert_allclose(q[np.isfinite(ref)], ref[np.isfinite(ref)], rtol=1e-5, atol=1e-5)
        best = pe.gather_best_q(case["party"], G["opponents"])
        np.testing.assert_allclose(best, np.asarray(case["best_q"], dtype=np.float32), rtol=1e-5, atol=1e-5)
        if "best_q_quick" in case:
            np.testing.assert_allclose(best, np.asarray(case["best_q_quick"], dtype=np.float32), rtol=1e-5, atol=1e-5)
    many = pe.gather_best_q_many(parties, G["opponents"])  # every party x opponent in one launch
    ref = np.asarray([c["best_q"] for c in G["cases"]], dtype=np.float32)
    np.testing.assert_allclose(many, ref, rtol=1e-5, atol=1e-5)
