"""cProfile of the blocking user call AgentVal.act_batch on pinned packed records (bench.py's `sync_value` path)."""
import cProfile
import pstats
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np
import torch

from pokeai_b200.agent import AgentVal
from pokeai_b200.feature_extractor import FeatureExtractor
from pokeai_b200.mlp_model import MLPModel
from pokeai_b200.packed import synth_packed
from pokeai_b200.session import pinned_records

B = 4096
fe = FeatureExtractor(6)
torch.manual_seed(0)
model = MLPModel(fe.input_shape, fe.output_dim).eval()
agent = AgentVal(model, fe)
rec = pinned_records(B)
rec[:] = synth_packed(B, seed=0)
for _ in range(20):
    agent.act_batch(rec)
t0 = time.perf_counter()
for _ in range(300):
    agent.act_batch(rec)
print("us per call", (time.perf_counter() - t0) / 300 * 1e6)
pr = cProfile.Profile()
pr.enable()
for _ in range(300):
    agent.act_batch(rec)
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(28)
