"""Tiny driver for ncu: three 32768-transition update steps (pk_dqn_grad + pk_clamp_adam) on one GPU."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np
import torch

from oracle import ref_port
from pokeai_b200 import ops
from pokeai_b200.packed import synth_packed
from pokeai_b200.tables import load_tables

dev = torch.device("cuda:0")
T = load_tables()
nl, H, A, n = 2, 64, 9, 32768
sd = ref_port.init_mlp_params(T.n_channels, nl, H, seed=1)
prep, _, _, flat = ops.prepare_weights(sd, dev)
prep_t = ops.prepare_weights(ref_port.init_mlp_params(T.n_channels, nl, H, seed=2), dev)[0]
rng = np.random.default_rng(0)
rs = synth_packed(n, seed=3, party_size=6, kind="full", tables=T)
rn = synth_packed(n, seed=4, party_size=6, kind="full", tables=T)
batch = [ops.packed_to_device(rs, dev), ops.packed_to_device(rn, dev),
         torch.from_numpy((rng.random(n) * rs["n_act"]).astype(np.int32)).to(dev),
         torch.from_numpy(rng.uniform(-1, 1, n).astype(np.float32)).to(dev),
         torch.from_numpy((rng.random(n) < 0.05).astype(np.uint8)).to(dev)]
m, v = torch.zeros_like(flat), torch.zeros_like(flat)
for step in (1, 2, 3):
    g, l = ops.dqn_grad(prep, prep_t, nl, H, A, *batch, gamma=0.95)
    ops.clamp_adam(flat, g, m, v, nl, H, step, prepared=prep)
torch.cuda.synchronize()
print("ok", float(l.item()))
