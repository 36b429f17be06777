"""Is the back-to-back launch rate of pk_decide bound by the GPU or by the host launch path?  Times B-state launches
(rotating output buffers) from a Python loop and replayed from a CUDA graph (no host work per launch)."""
import os, sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np, torch
from oracle import ref_port
from pokeai_b200 import ops
from pokeai_b200.packed import synth_packed
from pokeai_b200.tables import load_tables
B = int(os.environ.get("B", 4096)); A = 9
dev = torch.device("cuda:0"); T = load_tables()
pk = ops.packed_to_device(synth_packed(B, seed=0, party_size=6, kind="full", tables=T), dev)
prepared, nl, H, _ = ops.prepare_weights(ref_port.init_mlp_params(T.n_channels, 2, 64, seed=0), dev)
n_rot = max(3, int(np.ceil(300e6 / (B * 617 * A * 4))))
feats = [torch.empty((B, 617, A), dtype=torch.float32, device=dev) for _ in range(n_rot)]
mask = torch.empty((B, A), dtype=torch.int32, device=dev); q = torch.empty((B, A), dtype=torch.float32, device=dev)
act = torch.empty((B,), dtype=torch.int32, device=dev)
def launch(i): ops.decide(pk, A, prepared, nl, H, out_feat=feats[i % n_rot], out_mask=mask, out_q=q, out_action=act)
for i in range(10): launch(i)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
n = 400
t0 = time.perf_counter(); e0.record()
for i in range(n): launch(i)
e1.record(); t_issue = time.perf_counter() - t0; torch.cuda.synchronize()
print(f"B={B} python loop: {e0.elapsed_time(e1) / n * 1e3:.2f} us per launch on the GPU timeline; host issue {t_issue / n * 1e6:.2f} us per launch")
st = torch.cuda.Stream()
with torch.cuda.stream(st):
    for i in range(3): launch(i)
    st.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=st):
        for i in range(n_rot * 4): launch(i)
    g.replay(); st.synchronize()
    e0.record(st)
    reps = 20
    for _ in range(reps): g.replay()
    e1.record(st); st.synchronize()
print(f"B={B} CUDA graph of {n_rot * 4} launches: {e0.elapsed_time(e1) / (reps * n_rot * 4) * 1e3:.2f} us per launch")
