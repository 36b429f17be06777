"""Tiny driver for ncu: materialise features once, then launches of the dense-input forward (pk_mlp_forward)."""
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch

from oracle import ref_port
from pokeai_b200 import ops
from pokeai_b200.packed import synth_packed
from pokeai_b200.tables import load_tables

B = int(os.environ.get("B", 65536))
dev = torch.device("cuda:0")
T = load_tables()
A = 9
rec = synth_packed(B, seed=0, party_size=6, kind="full", tables=T)
pk = ops.packed_to_device(rec, dev)
sd = ref_port.init_mlp_params(T.n_channels, 2, 64, seed=0)
prepared, nl, H, _ = ops.prepare_weights(sd, dev)
feat, mask = ops.encode(pk, A)
q = torch.empty((B, A), dtype=torch.float32, device=dev)
for rep in range(3):
    ops.mlp_forward(prepared, nl, H, feat, out=q)
torch.cuda.synchronize()
print("ok", float(q.sum()))
