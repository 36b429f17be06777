"""Scratch timing of the individual entry points on one GPU (CUDA events, rotating output buffers
larger than L2).  Not the benchmark -- bench.py is; this is for kernel iteration."""
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np
import torch

from oracle import ref_port
from pokeai_b200 import ops
from pokeai_b200.packed import synth_packed
from pokeai_b200.tables import load_tables


def timeit(fn, n_rot, iters=40, warm=5):
    for i in range(warm):
        fn(i % n_rot)
    torch.cuda.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for i in range(iters):
        evs[i][0].record()
        fn(i % n_rot)
        evs[i][1].record()
    torch.cuda.synchronize()
    ts = sorted(a.elapsed_time(b) * 1e3 for a, b in evs)
    s0 = torch.cuda.Event(enable_timing=True); s1 = torch.cuda.Event(enable_timing=True)
    s0.record()
    for i in range(iters):
        fn(i % n_rot)
    s1.record()
    torch.cuda.synchronize()
    return ts[len(ts) // 2], ts[0], s0.elapsed_time(s1) * 1e3 / iters


def main():
    B = int(os.environ.get("B", 4096))
    nl, H = int(os.environ.get("NL", 2)), int(os.environ.get("H", 64))
    dev = torch.device("cuda:0")
    T = load_tables()
    A = 9
    rec = synth_packed(B, seed=0, party_size=6, kind="full", tables=T)
    pk = ops.packed_to_device(rec, dev)
    sd = ref_port.init_mlp_params(T.n_channels, nl, H, seed=0)
    prepared, _, _, _ = ops.prepare_weights(sd, dev)
    bytes_feat = B * 617 * A * 4
    n_rot = max(2, int(np.ceil(400e6 / bytes_feat)))
    feats = [torch.empty((B, 617, A), dtype=torch.float32, device=dev) for _ in range(n_rot)]
    mask = torch.empty((B, A), dtype=torch.int32, device=dev)
    q = torch.empty((B, A), dtype=torch.float32, device=dev)
    act = torch.empty((B,), dtype=torch.int32, device=dev)
    alg = B * 22432
    rows = []
    rows.append(("encode", timeit(lambda i: ops.encode(pk, A, out_feat=feats[i], out_mask=mask), n_rot)))
    rows.append(("decide+feat", timeit(lambda i: ops.decide(pk, A, prepared, nl, H, out_feat=feats[i], out_mask=mask, out_q=q, out_action=act), n_rot)))
    rows.append(("decide nofeat", timeit(lambda i: ops.decide(pk, A, prepared, nl, H, materialise=False, out_mask=mask, out_q=q, out_action=act), n_rot)))
    rows.append(("dense fwd", timeit(lambda i: ops.mlp_forward(prepared, nl, H, feats[i], out=q), n_rot)))
    big = torch.empty(int(400e6) // 4, dtype=torch.float32, device=dev)
    rows.append(("memset 400MB/n_rot", tuple(x / n_rot for x in timeit(lambda i: big.zero_(), 1))))
    print(f"B={B} net={nl}x{H} n_rot={n_rot} env={ {k: v for k, v in os.environ.items() if k.startswith('PK_')} }")
    for name, (med, mn, b2b) in rows:
        print(f"{name:20s} median {med:8.1f} us  min {mn:8.1f} us  back-to-back {b2b:8.1f} us   "
              f"{alg / (b2b * 1e-6) / 1e9:8.1f} GB/s(alg, b2b)  {B / (b2b * 1e-6) / 1e6:8.2f} Mdec/s")


if __name__ == "__main__":
    main()
