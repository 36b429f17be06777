"""Diagnostic: train-mode BatchNorm forward / backward against a float64 torch statement, per parameter tensor."""
import copy
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np
import torch
import torch.nn.functional as F

from pokeai_b200 import bn_train, ops
from pokeai_b200.mlp_model import MLPModel
from pokeai_b200.packed import synth_packed
from pokeai_b200.tables import load_tables

T = load_tables()
dev = torch.device("cuda:0")
ops.N.ensure_init(0)


def fwd(model, x):
    h = x
    acts = []
    for conv, bn in zip(model.layers, model.bn_layers):
        h = F.relu(bn(conv(h)))
        acts.append(h)
    return model.output(h).view(x.shape[0], -1), acts


for nl, H, B, ps in [(2, 64, 700, 6), (3, 96, 257, 6), (3, 64, 257, 6), (2, 96, 257, 6)]:
    A = 4 + ps - 1
    torch.manual_seed(nl * 100 + H)
    model = MLPModel((T.n_channels, A), A, n_layers=nl, n_channels=H, bn=True)
    with torch.no_grad():
        for b in model.bn_layers:
            b.weight.uniform_(0.5, 1.5)
            b.bias.normal_(0, 0.1)
    feat, _ = ops.encode(ops.packed_to_device(synth_packed(B, seed=3, party_size=ps, kind="ragged", tables=T), dev), A)
    m64 = copy.deepcopy(model).double()
    for p in m64.parameters():
        p.requires_grad_(True)
    q64, acts = fwd(m64, feat.cpu().double())
    g = torch.Generator().manual_seed(1)
    dq = torch.randn(B, A, generator=g) / B
    (q64 * dq.double()).sum().backward()
    flat = bn_train.flat_parameters(model, dev)
    q, ws = bn_train.forward_train(flat, None, nl, H, feat)
    grad = torch.empty_like(flat)
    bn_train.backward(flat, nl, H, feat, ws, dq.to(dev), grad)
    print(f"{nl}x{H} B={B}: q err {float((q.cpu().double() - q64.detach()).abs().max()):.2e}")
    near = [float(a.detach()[a.detach() > 0].min()) for a in acts]
    print("   smallest positive post-ReLU activation per layer (fp64):", ["%.2e" % v for v in near])
    off = 0
    for k, p in m64.named_parameters():
        n = p.numel()
        d = (grad[off:off + n].cpu().double() - p.grad.reshape(-1)).abs()
        print(f"   {k:28s} max|g| {float(p.grad.abs().max()):.3e}  err {float(d.max()):.2e}  at {int(d.argmax())}")
        off += n
