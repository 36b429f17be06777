"""Tiny driver for ncu: wide network 4 x 2048 at B states, forward (and backward with TRAIN=1), a few rounds."""
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch

from oracle import ref_port
from pokeai_b200 import ops, wide
from pokeai_b200.packed import synth_packed
from pokeai_b200.tables import load_tables

B = int(os.environ.get("B", 4096))
TRAIN = int(os.environ.get("TRAIN", 0))
dev = torch.device("cuda:0")
T = load_tables()
A, nl, H = 9, 4, 2048
pk = ops.packed_to_device(synth_packed(B, seed=0, party_size=6, kind="full", tables=T), dev)
prepared, _, _ = wide.prepare_wide(ref_port.init_mlp_params(T.n_channels, nl, H, seed=0), dev)
if TRAIN:
    ws = wide.train_workspace(nl, H, B, A, dev)
    dq = (torch.randn((B, A)) / (B * A)).to(dev)
    for _ in range(2):
        wide.train_forward(pk, A, prepared, nl, H, ws)
        g = wide.train_backward(prepared, nl, H, B, A, ws, dq)
    torch.cuda.synchronize()
    print("ok", float(g.abs().sum()))
else:
    ws = wide.workspace(nl, H, B, A, dev)
    for _ in range(3):
        m, q, a = wide.wide_forward(pk, A, prepared, nl, H, ws=ws)
    torch.cuda.synchronize()
    print("ok", int(a.sum()))
