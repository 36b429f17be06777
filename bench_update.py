#!/usr/bin/env python
"""bench_update.py -- transitions/sec of the DQN minibatch update (BASELINE.json configs[3]).

    python bench_update.py [--gpus N] [--steps K] [--warmup W] [--batch 32768]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench_update.py --gpus N --steps K --warmup W

One "step" = one optimiser step of `Trainer._update` (reference trainer.py:152-195) on a GLOBAL minibatch of
`--batch` packed transitions: online + target forward over next_state (Double DQN), forward/backward of the
chosen action column, fixed-order gradient reduction, [NCCL sum-allreduce of the flat gradient when N > 1],
clamp, Adam, refresh of the prepared weights.  The global batch is fixed and each rank takes a contiguous
shard => strong scaling.  `value` is measured with the shard resident in HBM (pk_dqn_grad + allreduce +
pk_clamp_adam, CUDA events, max over ranks); `e2e` goes through `Trainer._update_from_arrays` with host numpy
arrays (copy into pinned staging, H2D, kernels, allreduce, D2H of the loss, every step).
`cpu_baseline` (rank 0, N = 1) is the oracle port of the reference update (torch CPU autograd on dense
features, oracle/ref_port.py) on a bounded sample.

The update works on packed records (2 x 144 B per transition), so SURVEY.md 8d's dense HBM figure does not
apply; the roofline entry reports both the algorithmic-bytes view and the fp32-FMA view and says which binds.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

import bench as B0  # ClockSampler, peaks, host_threads

METRIC = "transitions/sec (DQN minibatch update: 2 next-state forwards + forward/backward + clamp + Adam)"
UNIT = "transitions/s"
A = 9
PARTY_SIZE = 6


def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=100)
    p.add_argument("--warmup", type=int, default=10)
    p.add_argument("--batch", type=int, default=32768, help="GLOBAL transitions per update")
    p.add_argument("--layers", type=int, default=2)
    p.add_argument("--channels", type=int, default=64)
    p.add_argument("--cpu-seconds", type=float, default=10.0)
    p.add_argument("--allreduce", default="peer", choices=["peer", "nccl"],
                   help="N > 1, default net: gradient exchange fused with clamp + Adam over NVLink peer memory "
                        "(pk_peer_clamp_adam), or NCCL allreduce + pk_clamp_adam")
    return p.parse_args()


def synth_transitions(n, seed):
    """5 % terminal, actions uniform over legal slots, rewards U(-1, 1)."""
    import numpy as np
    from pokeai_b200.packed import synth_packed
    rng = np.random.default_rng(seed)
    s = synth_packed(n, seed=seed, party_size=PARTY_SIZE, kind="full")
    s2 = synth_packed(n, seed=seed + 1, party_size=PARTY_SIZE, kind="ragged")
    n_act = s["n_act"].astype(np.int64)
    action = (rng.random(n) * np.maximum(n_act, 1)).astype(np.int32)
    reward = rng.uniform(-1, 1, n).astype(np.float32)
    terminal = (rng.random(n) < 0.05).astype(np.uint8)
    return s, s2, action, reward, terminal


def cpu_update_baseline(n_sample, nl, H, budget_s):
    """oracle port of Trainer._update on dense features (what the reference does), torch CPU, all host threads"""
    import numpy as np
    import torch
    from oracle import packed_oracle, ref_port
    from pokeai_b200.tables import load_tables
    T = load_tables()
    threads = B0.host_threads()
    torch.set_num_threads(threads)
    s, s2, a, r, t = synth_transitions(n_sample, 99)
    f, _ = packed_oracle.encode(s, PARTY_SIZE, T)
    f2, m2 = packed_oracle.encode(s2, PARTY_SIZE, T)
    sd = ref_port.init_mlp_params(T.n_channels, nl, H, seed=0)
    tsd = {k: v.clone() for k, v in sd.items()}
    times = []
    t_end = time.perf_counter() + budget_s
    while time.perf_counter() < t_end or len(times) < 2:
        t0 = time.perf_counter()
        ref_port.dqn_update_grads(sd, tsd, f, a.astype(np.int64), r, f2, m2, t == 0, 0.95, True)
        times.append(time.perf_counter() - t0)
        if len(times) >= 20:
            break
    med = sorted(times)[len(times) // 2]
    return {"value": n_sample / med, "unit": UNIT, "cores": threads, "kind": "port",
            "sample": f"{n_sample} dense transitions per update (features pre-encoded, as the reference's replay holds "
                      f"them), {len(times)} updates, median; loss + autograd backward, Adam not included"}


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", 0))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    import numpy as np
    import torch
    import torch.distributed as dist
    from pokeai_b200 import ops
    from pokeai_b200.dist import shard_bounds
    from pokeai_b200.trainer import Trainer

    if not torch.cuda.is_available():
        raise SystemExit("bench_update.py needs a CUDA device (pokeai_b200 has no CPU path)")
    torch.cuda.set_device(local_rank)
    dev = torch.device(f"cuda:{local_rank}")
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    NG, nl, H = args.batch, args.layers, args.channels
    lo, hi = shard_bounds(NG, world, rank)
    n = hi - lo
    torch.manual_seed(0)
    tr = Trainer({"n_layers": nl, "n_channels": H, "bn": False}, {"batch_size": NG, "lr": 1e-4},
                 {"party_size": PARTY_SIZE}, device=dev)
    n_rot = 4
    host_sets = []
    dev_sets = []
    for i in range(n_rot):  # every rank generates the same global minibatch and keeps its shard
        s, s2, a, r, t = synth_transitions(NG, 10 + i)
        hs = tuple(np.ascontiguousarray(x[lo:hi]) for x in (s, s2, a, r, t))
        host_sets.append(hs)
        dev_sets.append((ops.packed_to_device(hs[0], dev), ops.packed_to_device(hs[1], dev),
                         torch.from_numpy(hs[2]).to(dev), torch.from_numpy(hs[3]).to(dev),
                         torch.from_numpy(hs[4]).to(dev)))
    Cn = ops.n_channels()
    from pokeai_b200 import _native as N
    wide_net = tr._wide
    scratch = None if wide_net else torch.empty(N.lib().pk_dqn_scratch_floats(nl, H, Cn, max(n, 1)),
                                                dtype=torch.float32, device=dev)
    opt = tr.optimizer
    step_no = [0]
    px = None
    if world > 1 and not wide_net and args.allreduce == "peer":
        from pokeai_b200.dist import PeerGradExchange
        px = PeerGradExchange.create(tr._flat.numel(), dev)  # None on every rank if the box cannot map peer memory

    def step(i):
        s, s2, a, r, t = dev_sets[i % n_rot]
        if wide_net:  # tensor-core path: 2 next-state forwards, chosen-column forward + backward, clamp + Adam
            overlap = world > 1 and args.allreduce != "nccl"  # per-layer buckets started from inside the backward
            loss = tr._wide_grad_device(s, s2, a, r, t, NG, dist if overlap else None)
            if world > 1 and not overlap:
                dist.all_reduce(tr._grad, op=dist.ReduceOp.SUM)
            tr._wide_apply()
            return loss
        if px is not None:  # shard gradient straight into the peer-mapped slot; ONE kernel: sum over ranks + clamp + Adam
            gslot, lslot = px.next_slot()
            ops.dqn_grad(tr._prepared, tr._prepared_target, nl, H, A, s, s2, a, r, t, tr.gamma, True, n_global=NG,
                         scratch=scratch, out_grad=gslot, out_loss=lslot)
            step_no[0] += 1
            px.reduce_clamp_adam(tr._flat, opt.exp_avg, opt.exp_avg_sq, nl, H, Cn, step_no[0], opt.lr, opt.betas, opt.eps,
                                 prepared=tr._prepared)
            return px.loss_status[:1]
        grad, loss = ops.dqn_grad(tr._prepared, tr._prepared_target, nl, H, A, s, s2, a, r, t, tr.gamma, True,
                                  n_global=NG, scratch=scratch)
        if world > 1:
            dist.all_reduce(grad, op=dist.ReduceOp.SUM)
        step_no[0] += 1
        ops.clamp_adam(tr._flat, grad, opt.exp_avg, opt.exp_avg_sq, nl, H, step_no[0], lr=opt.lr, betas=opt.betas,
                       eps=opt.eps, prepared=tr._prepared)
        return loss

    sampler = B0.ClockSampler(local_rank)
    sampler.start()
    for i in range(max(3, args.warmup)):
        step(i)
    torch.cuda.synchronize()
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    ev0.record()
    for i in range(args.steps):
        loss = step(i)
    ev1.record()
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    barrier()
    dev_ms = max_over_ranks(ev0.elapsed_time(ev1))
    clocks = sampler.summary(t0, t1)
    last_loss = float(loss.item())

    # ---- end to end: host arrays -> Trainer._update_from_arrays (staging copy, H2D, kernels, allreduce, loss D2H)
    for i in range(3):
        tr._update_from_arrays(*host_sets[i % n_rot], n_global=NG)
    torch.cuda.synchronize()
    barrier()
    te0 = time.perf_counter()
    for i in range(args.steps):
        tr._update_from_arrays(*host_sets[i % n_rot], n_global=NG)
    torch.cuda.synchronize()
    te = max_over_ranks(time.perf_counter() - te0)
    barrier()
    sampler.stop()
    e2e = {"value": NG * args.steps / te, "unit": UNIT, "h2d_bytes_per_step": n * (2 * 144 + 9),
           "d2h_bytes_per_step": 4, "ms_per_step": te / args.steps * 1e3,
           "api": "Trainer._update_from_arrays(host numpy shard): pinned staging copy, H2D, pk_dqn_grad, "
                  "[NCCL allreduce], pk_clamp_adam, loss.item()"}
    cpu = None
    if rank == 0 and world == 1 and args.cpu_seconds > 0:
        cpu = cpu_update_baseline(2048, nl, H, args.cpu_seconds)
    if rank == 0:
        ms = dev_ms / args.steps
        peak, peak_src = B0.peaks()
        alg_bytes = n * (2 * 144 + 9)  # packed state + next_state + action/reward/terminal per transition
        # useful fp32 FMAs per transition in gather form: two 9-column forwards of next_state, one column
        # forward + backward of the chosen action (weight gradients of the 64x64 layer dominate)
        hid = (nl - 1) * H * H
        fma_per = 2 * (50 * H + 9 * (8 * H + hid + H)) + (58 * H + hid + H) + (2 * hid + 58 * H + 2 * H)
        if wide_net:  # dense GEMMs: 2 x 9-column forwards + 1-column forward, dW of every layer, dX of all but layer 0
            fwd = H * (617 + (nl - 1) * H + 1)
            fma_per = 2 * 9 * fwd + fwd + fwd + (nl - 1) * H * H
        line = {
            "metric": METRIC, "value": NG * args.steps / (dev_ms * 1e-3), "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": ms, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": ("bf16 (fp32 accumulate)" if wide_net else "f32"), "data": "synthetic",
            "config": {"workload": f"DQN minibatch update, {NG} packed transitions per step (BASELINE configs[3]), "
                                   f"default MLP {nl}x{H}, Double DQN, gamma 0.95, 5% terminal",
                       "global_transitions_per_step": NG, "transitions_per_gpu": n,
                       "sharding": f"dp{world}: contiguous shards of one global minibatch; " +
                                   ("gradient sum over ranks + clamp + Adam in ONE kernel over NVLink peer memory "
                                    f"(pk_peer_clamp_adam, {tr._flat.numel()} fp32 gradients, rank-ordered sum)"
                                    if px is not None else
                                    f"one NCCL sum-allreduce of {tr._flat.numel()} fp32 gradients per step, clamp + Adam "
                                    "after the reduce"),
                       "exchange": "peer" if px is not None else
                                   ("none" if world == 1 else
                                    "nccl, one bucket per layer overlapped with the backward"
                                    if wide_net and args.allreduce != "nccl" else "nccl"),
                       "l2": f"{n_rot} rotating transition sets; inputs are 297 B per transition (9.7 MB per set), "
                             "the kernels are compute/latency-bound, not HBM-bound"},
            "roofline": {"bound": "tensor (bf16 tcgen05 GEMMs)" if wide_net else
                         "fp32-fma (packed replay moved the update off the HBM roofline)",
                         "achieved": alg_bytes / (ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                         "frac": alg_bytes / (ms * 1e-3) / 1e9 / peak, "traffic": None,
                         "algorithmic_bytes_per_step": alg_bytes, "peak_source": peak_src,
                         "fma_per_transition": fma_per,
                         "achieved_tflops": 2 * fma_per * n / (ms * 1e-3) / 1e12,
                         "peak_tflops_fp32_nominal": 148 * 128 * 2 * 1.965e9 / 1e12},
            "e2e": e2e, "cpu_baseline": cpu, "clocks": clocks, "gpu_launches_per_step": (3 * (nl + 3) + 5 * nl + 8) if wide_net else 7,
            "last_loss": last_loss,
        }
        if wide_net:  # tensor roofline: useful FLOPs / time against the measured sustained bf16 rate
            import bench_wide
            tpeak, tsrc = bench_wide.tensor_peak()
            ach = 2 * fma_per * n / (ms * 1e-3) / 1e12
            line["roofline"] = {"bound": "tensor", "achieved": ach, "peak": tpeak, "unit": "TFLOP/s", "frac": ach / tpeak,
                                "traffic": None, "kernel": "pk_wide_gemm_kernel", "flops_per_transition": 2 * fma_per,
                                "peak_source": tsrc}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
