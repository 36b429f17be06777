"""Timeline of pk_dense_tc_kernel's three roles in CTA 0 (debug trace, pk_debug_set_trace): SM cycles between the
trace points of successive k-block iterations.  usage: B=65536 python tools/trace_dense.py"""
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np
import torch

from oracle import ref_port
from pokeai_b200 import _native as N, ops
from pokeai_b200.packed import synth_packed
from pokeai_b200.tables import load_tables

B = int(os.environ.get("B", 65536))
dev = torch.device("cuda:0")
T = load_tables()
A = 9
pk = ops.packed_to_device(synth_packed(B, seed=0, party_size=6, kind="full", tables=T), dev)
sd = ref_port.init_mlp_params(T.n_channels, 2, 64, seed=0)
prepared, nl, H, _ = ops.prepare_weights(sd, dev)
feat, _ = ops.encode(pk, A)
q = torch.empty((B, A), dtype=torch.float32, device=dev)
for _ in range(3):
    ops.mlp_forward(prepared, nl, H, feat, out=q)
tr = torch.zeros((3, 128, 8), dtype=torch.int64, device=dev)
N.check(N.lib().pk_debug_set_trace(tr.data_ptr()))
ops.mlp_forward(prepared, nl, H, feat, out=q)
torch.cuda.synchronize()
N.check(N.lib().pk_debug_set_trace(None))
t = tr.cpu().numpy()
t0 = t[t > 0].min()
roles = {0: ["top", "x_full", "lds issued", "a_empty", "tmem st", "arrived", "(tile end) loop out", "(tile end) head done"],
         1: ["top", "a_full", "w_full", "issued", "committed"],
         2: ["W top", "w_empty", "X top", "x_empty", "x issued"]}
lo, hi = int(os.environ.get("IT0", 20)), int(os.environ.get("IT1", 44))
for r, names in roles.items():
    print(f"role {r}: cycles since kernel start at each point, iterations {lo}..{hi - 1}")
    print("   it " + " ".join(f"{n:>10s}" for n in names))
    for it in range(lo, hi):
        row = [int(t[r, it, p] - t0) if t[r, it, p] else -1 for p in range(len(names))]
        print(f"  {it:3d} " + " ".join(f"{v:10d}" for v in row))
per = np.diff(t[1, 20:120, 3].astype(np.int64))
print("MMA issue period (cycles): median", np.median(per), "mean", per.mean())
