"""Multi-GPU check of the fused peer-memory gradient exchange (run under torchrun, one rank per GPU):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29540 \
        tools/check_peer_update.py

Both routes start from the same weights and run the same sharded updates: (a) pk_dqn_grad -> NCCL allreduce ->
pk_clamp_adam, (b) pk_dqn_grad into the peer buffer -> pk_peer_clamp_adam.  Checks: reduced gradient and loss equal to
NCCL's within fp32 summation-order noise; parameters after every step equal within 1e-6; the peer route's parameters
are bit-identical on every rank; the prepared copy equals a fresh preparation of the new parameters."""
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np
import torch
import torch.distributed as dist

from oracle import ref_port
from pokeai_b200 import ops
from pokeai_b200.dist import PeerGradExchange, shard_bounds
from pokeai_b200.packed import synth_packed
from pokeai_b200.tables import load_tables


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device(f"cuda:{int(os.environ.get('LOCAL_RANK', rank))}")
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    T = load_tables()
    nl, H, A, n = 2, 64, 9, 4096
    Cn = T.n_channels
    sd = ref_port.init_mlp_params(Cn, nl, H, seed=1)
    sd_t = ref_port.init_mlp_params(Cn, nl, H, seed=2)
    prep_a, _, _, flat_a = ops.prepare_weights(sd, dev)
    prep_b, _, _, flat_b = ops.prepare_weights(sd, dev)
    prep_t = ops.prepare_weights(sd_t, dev)[0]
    n_par = flat_a.numel()
    px = PeerGradExchange.create(n_par, dev)
    assert px is not None, "peer exchange unavailable on this box"
    ma, va = torch.zeros_like(flat_a), torch.zeros_like(flat_a)
    mb, vb = torch.zeros_like(flat_b), torch.zeros_like(flat_b)
    gsum = torch.empty_like(flat_b)
    lo, hi = shard_bounds(n, world, rank)
    worst = 0.0
    for step in range(1, 7):
        rng = np.random.default_rng(100 + step)
        rs = synth_packed(n, seed=10 + step, party_size=6, kind="ragged", tables=T)
        rn = synth_packed(n, seed=50 + step, party_size=6, kind="ragged", tables=T)
        act = (rng.random(n) * rs["n_act"]).astype(np.int32)
        rew = rng.uniform(-2, 2, n).astype(np.float32)
        term = (rng.random(n) < 0.05)
        batch = [ops.packed_to_device(rs[lo:hi], dev), ops.packed_to_device(rn[lo:hi], dev),
                 torch.from_numpy(act[lo:hi]).to(dev), torch.from_numpy(rew[lo:hi]).to(dev),
                 torch.from_numpy(term[lo:hi].astype(np.uint8)).to(dev)]
        # (a) NCCL route
        g, l = ops.dqn_grad(prep_a, prep_t, nl, H, A, *batch, gamma=0.95, n_global=n)
        dist.all_reduce(g)
        dist.all_reduce(l)
        ops.clamp_adam(flat_a, g, ma, va, nl, H, step, prepared=prep_a)
        # (b) peer route: same gradient kernels, written into this step's peer slot
        gslot, lslot = px.next_slot()
        g2, l2 = ops.dqn_grad(prep_b, prep_t, nl, H, A, *batch, gamma=0.95, n_global=n)
        gslot.copy_(g2)
        lslot.copy_(l2)
        px.reduce_clamp_adam(flat_b, mb, vb, nl, H, Cn, step, 1e-3, (0.9, 0.999), 1e-8, prepared=prep_b, grad_sum=gsum)
        torch.cuda.synchronize()
        assert float(px.loss_status[1]) == 0.0, "a peer did not arrive"
        np.testing.assert_allclose(gsum.cpu().numpy(), g.cpu().numpy(), rtol=1e-5, atol=1e-8)
        assert abs(float(px.loss_status[0]) - float(l.item())) < 1e-6
        d = float((flat_a - flat_b).abs().max())
        worst = max(worst, d)
        assert d < 2e-6, (step, d)
        assert torch.equal(ops.prepare_from_flat(flat_b, nl, H, Cn), prep_b)
        allp = [torch.empty_like(flat_b) for _ in range(world)]
        dist.all_gather(allp, flat_b)
        assert all(torch.equal(allp[0], p) for p in allp), "peer route: parameters differ between ranks"
    dist.barrier()
    if rank == 0:
        print(f"peer update ok: world {world}, 6 steps, max |param(NCCL route) - param(peer route)| = {worst:.2e}, "
              f"peer-route parameters bit-identical on all ranks")
    dist.destroy_process_group()


main()
