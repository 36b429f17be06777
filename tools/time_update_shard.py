"""Diagnostic: the per-rank part of a data-parallel update (pk_dqn_grad on a shard of a 4 096-transition minibatch +
pk_clamp_adam) timed on ONE GPU for shards of 4096 / 2048 / 1024 / 512 transitions, i.e. what each rank of 1 / 2 / 4 / 8
runs between exchanges.  Shows how far the step is from shrinking with the shard.  Not a bench line."""
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


def main():
    dev = torch.device("cuda:0")
    T = load_tables()
    nl, H, A, NG = 2, 64, 9, 4096
    sd = ref_port.init_mlp_params(T.n_channels, nl, H, seed=0)
    prepared, _, _, flat = ops.prepare_weights(sd, dev)
    prepared_t = ops.prepare_weights(ref_port.init_mlp_params(T.n_channels, nl, H, seed=1), dev)[0]
    rng = np.random.default_rng(0)
    rs = synth_packed(NG, seed=1, party_size=6, kind="ragged", tables=T)
    rn = synth_packed(NG, seed=2, party_size=6, kind="ragged", tables=T)
    act = (rng.random(NG) * rs["n_act"]).astype(np.int32)
    rew = rng.uniform(-1, 1, NG).astype(np.float32)
    ter = (rng.random(NG) < 0.05).astype(np.uint8)
    m, v = torch.zeros_like(flat), torch.zeros_like(flat)
    shards = [int(x) for x in os.environ.get("SHARDS", "4096,2048,1024,512").split(",")]
    iters = int(os.environ.get("ITERS", 200))
    for n in shards:
        s, s2 = ops.packed_to_device(rs[:n], dev), ops.packed_to_device(rn[:n], dev)
        a, r, t = (torch.from_numpy(x[:n]).to(dev) for x in (act, rew, ter))
        grad, loss = ops.dqn_grad(prepared, prepared_t, nl, H, A, s, s2, a, r, t, gamma=0.95, n_global=NG)
        scratch = torch.empty(ops.N.lib().pk_dqn_scratch_floats(nl, H, T.n_channels, n), dtype=torch.float32, device=dev)

        def step(k):
            ops.dqn_grad(prepared, prepared_t, nl, H, A, s, s2, a, r, t, gamma=0.95, n_global=NG, scratch=scratch,
                         out_grad=grad, out_loss=loss)
            ops.clamp_adam(flat, grad, m, v, nl, H, k + 1, prepared=prepared)

        for k in range(10):
            step(k)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for k in range(iters):
            step(k)
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / iters
        print(f"shard {n:5d} of {NG}: {us:7.1f} us per step (stream back-to-back, eager launches)  "
              f"ideal from 4096: -  per-transition {us / n * 1e3:7.1f} ns")


if __name__ == "__main__":
    main()
