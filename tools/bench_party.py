"""Throughput of the party evaluator: P parties x all 251 opponent species per call, host parties in,
best-Q matrix out (pack on the host + H2D + pk_decide without features + D2H).  The reference's hill-climb does this
one party at a time at batch ~129 (party_evaluator_quick.py:97-104).  usage: P=100 python tools/bench_party.py"""
import os
import random
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np
import torch

from pokeai_b200 import AgentVal, FeatureExtractor, MLPModel, PartyEvaluatorQuick
from pokeai_b200.tables import load_tables

P = int(os.environ.get("P", 100))
T = load_tables()
dev = torch.device("cuda:0")
fe = FeatureExtractor(6, device=dev)
torch.manual_seed(0)
model = MLPModel(fe.input_shape, fe.output_dim).eval()
pe = PartyEvaluatorQuick(AgentVal(model, fe), 6)
rng = random.Random(0)
species = list(T.species_id_to_idx)
moves = list(T.move_to_idx)
parties = [[{"species": sp, "moves": rng.sample(moves, 4), "item": ""} for sp in rng.sample(species, 6)] for _ in range(P)]
pe.gather_best_q_many(parties[:2], species)
t0 = time.perf_counter()
rec = np.concatenate([pe.pack(p, species) for p in parties])
t1 = time.perf_counter()
reps = 20
for _ in range(reps):
    best = pe._q(rec).max(axis=1)
t2 = time.perf_counter()
n = len(rec)
print(f"{P} parties x {len(species)} opponents = {n} decisions: host packing {1e3 * (t1 - t0):.1f} ms "
      f"({n / (t1 - t0) / 1e6:.2f} M/s), records -> best Q {1e3 * (t2 - t1) / reps:.2f} ms "
      f"({n * reps / (t2 - t1) / 1e6:.1f} M decisions/s end to end), checksum {float(best.sum()):.4f}")
