"""Scratch: per-launch time of the 4096-state decision launch when the 16 independent launches of a bench step are
captured on S streams (fork / join inside one CUDA graph) instead of one.  usage: python tools/quick_streams.py"""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np
import torch

from pokeai_b200 import _native as N, ops
from pokeai_b200.feature_extractor import FeatureExtractor
from pokeai_b200.mlp_model import MLPModel
from pokeai_b200.packed import synth_packed

dev = torch.device("cuda:0")
B, A, G = 4096, 9, 16
fe = FeatureExtractor(6, device=dev)
Cn = fe.input_shape[0]
torch.manual_seed(0)
model = MLPModel(fe.input_shape, fe.output_dim).eval()
prepared = model.prepared(dev)
n_rot = 4
pk = [ops.packed_to_device(synth_packed(B, seed=i, party_size=6, kind="full"), dev) for i in range(n_rot)]
feats = [torch.empty((B, Cn, A), dtype=torch.float32, device=dev) for _ in range(n_rot)]
masks = [torch.empty((B, A), dtype=torch.int32, device=dev) for _ in range(n_rot)]
qs = [torch.empty((B, A), dtype=torch.float32, device=dev) for _ in range(n_rot)]
acts = [torch.empty((B,), dtype=torch.int32, device=dev) for _ in range(n_rot)]


def step(i):
    j = i % n_rot
    ops.decide(pk[j], A, prepared, 2, 64, out_feat=feats[j], out_mask=masks[j], out_q=qs[j], out_action=acts[j])


for pdl in (1, 0):
    N.check(N.lib().pk_set_overlap_launches(pdl))
    for S in (1, 2, 4):
        main = torch.cuda.Stream(device=dev)
        others = [torch.cuda.Stream(device=dev) for _ in range(S - 1)]
        streams = [main] + others
        with torch.cuda.stream(main):
            for i in range(G):
                step(i)
            main.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=main):
                fork = torch.cuda.Event()
                fork.record(main)
                for s in others:
                    s.wait_event(fork)
                for i in range(G):
                    with torch.cuda.stream(streams[i % S]):
                        step(i)
                for s in others:
                    e = torch.cuda.Event()
                    e.record(s)
                    main.wait_event(e)
            for _ in range(5):
                graph.replay()
            main.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(main)
            for _ in range(20):
                graph.replay()
            e1.record(main)
            main.synchronize()
            us = e0.elapsed_time(e1) * 1e3 / (20 * G)
            print(f"pdl={pdl} streams={S}: {us:.2f} us per launch, {B / us:.1f} M decisions/s, frac {22432 * B / us / 1e3 / 6535.1:.3f}")
