"""Diagnostic: a few bn: true updates of 4 096 transitions (for an ncu launch list, or plain timing with host phases)."""
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np
import torch

import bench_secondary as bs
from pokeai_b200.trainer import Trainer

NG = 4096
torch.manual_seed(11)
tr = Trainer({"n_layers": 2, "n_channels": 64, "bn": True}, {"batch_size": NG, "lr": 1e-4}, {"party_size": 6}, device="cuda:0")
sets = [bs.synth_transitions(NG, 40 + i) for i in range(2)]
for i in range(3):
    tr._update_from_arrays(*sets[i % 2])
torch.cuda.synchronize()
t0 = time.perf_counter()
n = 10
for i in range(n):
    tr._update_from_arrays(*sets[i % 2])
torch.cuda.synchronize()
print(f"bn update, {NG} transitions: {(time.perf_counter() - t0) / n * 1e3:.3f} ms per step (wall)")
