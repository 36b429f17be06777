"""Phase timeline of the decision kernel (debug trace, pk_debug_set_trace): median over warps of the time spent
between the 10 trace points of each warp's first battle.  usage: B=4096 python tools/trace_phases.py"""
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np
import torch

from pokeai_b200 import _native as N, ops
from pokeai_b200.mlp_model import MLPModel
from pokeai_b200.packed import synth_packed

B = int(os.environ.get("B", 4096))
WARPS = int(os.environ.get("PK_DECIDE_WARPS", 16))
dev = torch.device("cuda:0")
A = 9
pk = ops.packed_to_device(synth_packed(B, seed=0), dev)
torch.manual_seed(0)
model = MLPModel((617, A), A).eval()
prepared = model.prepared(dev)
feat = torch.empty((B, 617, A), dtype=torch.float32, device=dev)
names = ["start->staging issued", "->first record parsed", "->prologue barrier", "->features issued", "->state part", "->gathers",
         "->hidden+head", "->select+stores", "->pokes done"]
for mat in (True, False):
    for _ in range(3):
        ops.decide(pk, A, prepared, 2, 64, materialise=mat, out_feat=feat if mat else None)
    n_slots = 148 * 32
    tr = torch.zeros((n_slots, 16), dtype=torch.int64, device=dev)
    N.check(N.lib().pk_debug_set_trace(tr.data_ptr()))
    ops.decide(pk, A, prepared, 2, 64, materialise=mat, out_feat=feat if mat else None)
    torch.cuda.synchronize()
    N.check(N.lib().pk_debug_set_trace(None))
    t = tr.cpu().numpy()
    t = t[(t[:, 0] > 0) & (t[:, 9] > 0)]
    t0 = t[:, 0].min()
    print(f"B={B} features={'yes' if mat else 'no'} warps traced={len(t)}; kernel span "
          f"{(t[:, 9].max() - t0) / 1e3:.1f} us (first battles only)")
    for i, nm in enumerate(names):
        d = (t[:, i + 1] - t[:, i]) / 1e3
        print(f"  {nm:24s} median {np.median(d):6.2f} us  p90 {np.percentile(d, 90):6.2f}  max {d.max():6.2f}")
    print(f"  first battle total       median {np.median((t[:, 9] - t[:, 1]) / 1e3):6.2f} us;  CTA start spread "
          f"{(t[:, 0].max() - t0) / 1e3:.2f} us")
