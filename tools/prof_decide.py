"""Tiny driver for ncu: warm up, then one launch each of encode / decide+features / decide without features."""
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch

from oracle import ref_port
from pokeai_b200 import ops
from pokeai_b200.packed import synth_packed
from pokeai_b200.tables import load_tables

B = int(os.environ.get("B", 4096))
dev = torch.device("cuda:0")
T = load_tables()
A = 9
rec = synth_packed(B, seed=0, party_size=6, kind="full", tables=T)
pk = ops.packed_to_device(rec, dev)
sd = ref_port.init_mlp_params(T.n_channels, 2, 64, seed=0)
prepared, nl, H, _ = ops.prepare_weights(sd, dev)
feat = torch.empty((B, 617, A), dtype=torch.float32, device=dev)
mask = torch.empty((B, A), dtype=torch.int32, device=dev)
q = torch.empty((B, A), dtype=torch.float32, device=dev)
act = torch.empty((B,), dtype=torch.int32, device=dev)
for rep in range(3):  # 2 warm-up rounds + the profiled one (ncu -s 6 -c 3 on pk_decide_kernel)
    ops.encode(pk, A, out_feat=feat, out_mask=mask)
    ops.decide(pk, A, prepared, nl, H, out_feat=feat, out_mask=mask, out_q=q, out_action=act)
    ops.decide(pk, A, prepared, nl, H, materialise=False, out_mask=mask, out_q=q, out_action=act)
torch.cuda.synchronize()
print("ok", int(act.sum()))
