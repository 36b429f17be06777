"""Multi-GPU check of the wide update's overlapped gradient exchange (run under torchrun, one rank per GPU):
per-layer NCCL buckets started from inside the backward (pk_wide_train_backward_ev) must give the same reduced
gradient as one allreduce after the backward."""
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np
import torch
import torch.distributed as dist

from pokeai_b200 import ops
from pokeai_b200.dist import shard_bounds
from pokeai_b200.packed import synth_packed
from pokeai_b200.tables import load_tables
from pokeai_b200.trainer import Trainer


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device(f"cuda:{int(os.environ.get('LOCAL_RANK', rank))}")
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    T = load_tables()
    n = 4096
    torch.manual_seed(0)
    tr = Trainer({"n_layers": 4, "n_channels": 2048, "bn": False}, {"batch_size": n, "lr": 1e-4}, {"party_size": 6},
                 device=dev)
    lo, hi = shard_bounds(n, world, rank)
    worst = 0.0
    for it in range(3):
        rng = np.random.default_rng(7 + it)
        rs = synth_packed(n, seed=20 + it, party_size=6, kind="ragged", tables=T)
        rn = synth_packed(n, seed=60 + it, party_size=6, kind="ragged", tables=T)
        a = torch.from_numpy((rng.random(n) * rs["n_act"]).astype(np.int32)[lo:hi]).to(dev)
        r = torch.from_numpy(rng.uniform(-1, 1, n).astype(np.float32)[lo:hi]).to(dev)
        t = torch.from_numpy((rng.random(n) < 0.05).astype(np.uint8)[lo:hi]).to(dev)
        s, s2 = ops.packed_to_device(rs[lo:hi], dev), ops.packed_to_device(rn[lo:hi], dev)
        tr._wide_grad_device(s, s2, a, r, t, n, None)
        dist.all_reduce(tr._grad)
        want = tr._grad.clone()
        tr._wide_grad_device(s, s2, a, r, t, n, dist)
        torch.cuda.synchronize()
        got = tr._grad
        scale = float(want.abs().max())
        d = float((got - want).abs().max())
        worst = max(worst, d / scale)
        assert d <= 1e-5 * scale, (it, d, scale)
    dist.barrier()
    if rank == 0:
        print(f"wide overlap ok: world {world}, 3 steps, max |bucketed - monolithic| / max|g| = {worst:.2e}")
    # peer-memory route (pk_peer_allreduce: sliced reduce-scatter / all-gather over NVLink, one kernel): the summed
    # gradient must equal NCCL's within summation order, and be BIT-identical on every rank
    px = tr._peer_exchange_wide(dist)
    if px is None:
        if rank == 0:
            print("wide peer allreduce skipped: peer memory not available")
    else:
        worst_p = 0.0
        for it in range(4):  # both buffer parities, twice
            rng = np.random.default_rng(70 + it)
            rs = synth_packed(n, seed=120 + it, party_size=6, kind="ragged", tables=T)
            rn = synth_packed(n, seed=160 + it, party_size=6, kind="ragged", tables=T)
            a = torch.from_numpy((rng.random(n) * rs["n_act"]).astype(np.int32)[lo:hi]).to(dev)
            r = torch.from_numpy(rng.uniform(-1, 1, n).astype(np.float32)[lo:hi]).to(dev)
            t = torch.from_numpy((rng.random(n) < 0.05).astype(np.uint8)[lo:hi]).to(dev)
            s, s2 = ops.packed_to_device(rs[lo:hi], dev), ops.packed_to_device(rn[lo:hi], dev)
            loss = tr._wide_grad_device(s, s2, a, r, t, n, None)
            want = tr._grad.clone()
            want_loss = loss.clone()
            dist.all_reduce(want)
            dist.all_reduce(want_loss)
            grad, loss_slot = px.next_slot()
            loss_slot.copy_(tr._wide_grad_device(s, s2, a, r, t, n, None, out_grad=grad))
            px.allreduce()
            torch.cuda.synchronize()
            assert float(px.loss_status[1].item()) == 0.0, "peer allreduce timed out"
            scale = float(want.abs().max())
            d = float((grad - want).abs().max())
            worst_p = max(worst_p, d / scale)
            assert d <= 1e-5 * scale, (it, d, scale)
            assert abs(float(loss_slot.item()) - float(want_loss.item())) < 1e-6
            gathered = [torch.empty_like(grad) for _ in range(world)]
            dist.all_gather(gathered, grad.contiguous())
            assert all(torch.equal(gathered[0], g) for g in gathered), "peer allreduce: ranks disagree"
        dist.barrier()
        if rank == 0:
            print(f"wide peer allreduce ok: world {world}, 4 steps, bit-identical across ranks, "
                  f"max |peer - nccl| / max|g| = {worst_p:.2e}")
        # Trainer._update_wide end to end by both routes from the same start (a second trainer with the peer route off):
        # same loss, parameters within summation order, identical on every rank; one rank gets an EMPTY shard once
        import os as _os
        _os.environ["PK_PEER_ALLREDUCE"] = "0"
        torch.manual_seed(0)
        tn = Trainer({"n_layers": 4, "n_channels": 2048, "bn": False}, {"batch_size": n, "lr": 1e-4}, {"party_size": 6},
                     device=dev)
        assert tn._peer_exchange_wide(dist) is None  # decided (and cached) while the switch is off: NCCL buckets
        _os.environ["PK_PEER_ALLREDUCE"] = "1"
        tn._flat.copy_(tr._flat)
        tn.optimizer.exp_avg.copy_(tr.optimizer.exp_avg)
        tn.optimizer.exp_avg_sq.copy_(tr.optimizer.exp_avg_sq)
        tn.optimizer.step_count = tr.optimizer.step_count
        tn._refresh_prepared()
        tr._refresh_prepared()
        for it in range(3):
            rng = np.random.default_rng(700 + it)
            rs = synth_packed(n, seed=220 + it, party_size=6, kind="ragged", tables=T)
            rn = synth_packed(n, seed=260 + it, party_size=6, kind="ragged", tables=T)
            a = (rng.random(n) * rs["n_act"]).astype(np.int32)
            r = rng.uniform(-1, 1, n).astype(np.float32)
            t = (rng.random(n) < 0.05).astype(np.uint8)
            l2, h2 = (lo, hi) if not (it == 1 and rank == world - 1) else (hi, hi)  # an empty shard on the last rank
            lp = tr._update_from_arrays(rs[l2:h2], rn[l2:h2], a[l2:h2], r[l2:h2], t[l2:h2], n_global=n)
            ln = tn._update_from_arrays(rs[l2:h2], rn[l2:h2], a[l2:h2], r[l2:h2], t[l2:h2], n_global=n)
            assert abs(lp - ln) < 1e-6, (it, lp, ln)
        assert tn._peer_exchange_wide(dist) is None and tr._peer_exchange_wide(dist) is not None
        dpar = float((tr._flat - tn._flat).abs().max())
        assert dpar <= 1e-6, dpar
        gathered = [torch.empty_like(tr._flat) for _ in range(world)]
        dist.all_gather(gathered, tr._flat)
        assert all(torch.equal(gathered[0], g) for g in gathered)
        dist.barrier()
        if rank == 0:
            print(f"wide trainer peer route ok: 3 updates (one with an empty shard), max |peer - nccl| parameter = {dpar:.2e}")
    dist.destroy_process_group()


main()
