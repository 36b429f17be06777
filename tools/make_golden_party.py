"""Golden vectors for the party evaluator, produced by the REAL reference classes
(pokeai/ai/party_feature/party_evaluator.py, party_evaluator_quick.py) in the build container.

    PYTHONDONTWRITEBYTECODE=1 python tools/make_golden_party.py

tests/golden/party_eval.json.gz: for party sizes 3 and 6, seeded parties and opponent lists with
  * FeatureExtractor features (sparse) and masks of PartyEvaluator._make_obs_vector(party, opponent),
  * (party_eval_weights.npz) MLPModel weights (2 x 64, torch.manual_seed) and PartyEvaluator.calc_q_func / gather_best_q outputs,
  * PartyEvaluatorQuick.gather_best_q (the reference's own fast path; equal to the slow one).
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

import numpy as np  # noqa: E402
import torch  # noqa: E402

from pokeai.ai.generic_move_model.agent import Agent  # noqa: E402
from pokeai.ai.generic_move_model.feature_extractor import FeatureExtractor  # noqa: E402
from pokeai.ai.generic_move_model.mlp_model import MLPModel  # noqa: E402
from pokeai.ai.party_feature.party_evaluator import PartyEvaluator  # noqa: E402
from pokeai.ai.party_feature.party_evaluator_quick import PartyEvaluatorQuick, sample_party  # noqa: E402
from pokeai.util import DATASET_DIR, json_load  # noqa: E402

pokes = json_load(DATASET_DIR.joinpath("all_pokemons.json"))
moves = json_load(DATASET_DIR.joinpath("all_moves_with_hiddenpower_type.json"))
items = json_load(DATASET_DIR.joinpath("all_items.json"))


def rand_party(rng, size):
    party = []
    for sp in rng.sample(pokes, size):
        party.append({"name": sp, "species": sp, "moves": rng.sample(moves, rng.randint(1, 4)), "ability": "No Ability",
                      "item": rng.choice(items), "level": 50, "gender": "M"})
    return party


out = {}
weights = {}
for size in (3, 6):
    rng = random.Random(100 + size)
    fe = FeatureExtractor(size)
    torch.manual_seed(size)
    model = MLPModel(fe.input_shape, fe.output_dim, n_layers=2, n_channels=64)
    agent = Agent(model, fe)
    pe = PartyEvaluator(agent, size)
    parties = ([[dict(p) for p in sample_party]] if size == 3 else []) + [rand_party(rng, size) for _ in range(4)]
    opponents = rng.sample(pokes, 12)
    cases = []
    for party in parties:
        feats, masks, qs = [], [], []
        for opp in opponents:
            f, m = pe._make_obs_vector(party, opp)
            idx = np.flatnonzero(f)
            feats.append({"idx": idx.tolist(), "val": [float(x) for x in f.ravel()[idx]]})
            masks.append([int(x) for x in m])
            qs.append([float(x) if np.isfinite(x) else None for x in pe.calc_q_func(party, opp)])
        best = pe.gather_best_q(party, opponents)
        cases.append({"party": party, "features": feats, "masks": masks, "q": qs, "best_q": [float(x) for x in best]})
    if size == 3:  # the reference's own fast path needs its 3-member sample party
        peq = PartyEvaluatorQuick(agent, size, "cpu")
        for c in cases:
            c["best_q_quick"] = [float(x) for x in peq.gather_best_q(c["party"], opponents)]
    out[str(size)] = {"opponents": opponents, "cases": cases, "shape": [int(fe.input_shape[0]), int(fe.input_shape[1])]}
    weights.update({f"{size}_{k}": v.detach().numpy() for k, v in model.state_dict().items()})
with gzip.open(ROOT / "tests" / "golden" / "party_eval.json.gz", "wt") as f:
    json.dump(out, f)
np.savez_compressed(ROOT / "tests" / "golden" / "party_eval_weights.npz", **weights)
print("written", (ROOT / "tests" / "golden" / "party_eval.json.gz").stat().st_size, "bytes")
