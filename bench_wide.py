#!/usr/bin/env python
"""bench_wide.py -- decisions/sec of the wide policy network (BASELINE.json configs[4]: 4 x 2048 hidden,
full species / move one-hot features), tensor-core bound.

    python bench_wide.py [--gpus N] [--steps K] [--warmup W] [--batch 4096] [--mode forward|train]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench_wide.py --gpus N ...

One "step" = one pass over `--batch` synthetic battle states per GPU (36 864 GEMM rows at 4 096 states):
  forward  packed records -> bf16 operand (B*A, 640) -> 4 fused GEMM+bias+ReLU layers on tcgen05 -> head -> masked
           argmax (pk_wide_forward).
  train    forward + backward (pk_wide_train_step when built) -- reported separately.
Battles are independent: every rank runs its own shard, no collective in the forward (weak scaling).
`roofline` is the TENSOR roofline: useful FLOPs (2*M*N*K per layer at the true K = 617 for layer 0) / CUDA-event
time against the measured dense bf16 cuBLAS throughput in MEASURED_PEAKS.json (sustained figure: the GEMMs run
back to back inside a long step).  `e2e` goes through WideRunner.decide with host records.
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

import bench as B0

METRIC = "policy decisions/sec, wide net 4x2048 (encode + tensor-core MLP forward + masked argmax)"
UNIT = "decisions/s"
A = 9
PARTY_SIZE = 6


def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=30)
    p.add_argument("--warmup", type=int, default=5)
    p.add_argument("--batch", type=int, default=4096)
    p.add_argument("--layers", type=int, default=4)
    p.add_argument("--channels", type=int, default=2048)
    p.add_argument("--cpu-seconds", type=float, default=10.0)
    p.add_argument("--mode", default="forward", choices=["forward", "train"],
                   help="forward: decisions (encode + forward + argmax); train: forward + backward to the flat gradient")
    return p.parse_args()


def N_params(nl, H, Cn):
    return H * Cn + H + (nl - 1) * (H * H + H) + H + 1


def tensor_peak():
    f = ROOT / "MEASURED_PEAKS.json"
    if f.exists():
        try:
            d = json.loads(f.read_text())
            return float(d["bf16_tflops_sustained"]), "measured (MEASURED_PEAKS.json bf16_tflops_sustained)"
        except Exception:
            pass
    return 1434.9, "fallback (B200_PROFILING.md sustained dense bf16)"


def cpu_baseline(nl, H, budget_s):
    """oracle port of the reference forward (torch CPU fp32 conv1d, all host threads) on a bounded sample"""
    import numpy as np
    import torch
    from oracle import packed_oracle, ref_port
    from pokeai_b200.packed import synth_packed
    from pokeai_b200.tables import load_tables
    T = load_tables()
    threads = B0.host_threads()
    torch.set_num_threads(threads)
    n = 256
    rec = synth_packed(n, seed=0, party_size=PARTY_SIZE, kind="full", tables=T)
    sd = ref_port.init_mlp_params(T.n_channels, nl, H, seed=0)
    times = []
    t_end = time.perf_counter() + budget_s
    while (time.perf_counter() < t_end or len(times) < 2) and len(times) < 20:
        t0 = time.perf_counter()
        f, m = packed_oracle.encode(rec, PARTY_SIZE, T)
        with torch.no_grad():
            q = ref_port.mlp_forward(sd, torch.from_numpy(f)).numpy()
        ref_port.masked_argmax(q, m)
        times.append(time.perf_counter() - t0)
    med = sorted(times)[len(times) // 2]
    return {"value": n / med, "unit": UNIT, "cores": threads, "kind": "port",
            "sample": f"{n} states per pass, {len(times)} passes, median; vectorised numpy encode + torch CPU fp32 "
                      f"forward ({threads} threads) + masked argmax"}


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", 0))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    import numpy as np
    import torch
    import torch.distributed as dist
    from pokeai_b200 import ops, wide
    from pokeai_b200.mlp_model import MLPModel
    from pokeai_b200.packed import synth_packed

    if not torch.cuda.is_available():
        raise SystemExit("bench_wide.py needs a CUDA device (pokeai_b200 has no CPU path)")
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

    B, nl, H = args.batch, args.layers, args.channels
    Cn = ops.n_channels()
    torch.manual_seed(0)
    model = MLPModel((Cn, A), A, n_layers=nl, n_channels=H).eval()
    prepared, _, _ = wide.prepare_wide(model.folded_state_dict(), dev)
    ws = wide.workspace(nl, H, B, A, dev)
    n_rot = 3
    recs = [synth_packed(B, seed=1000 * rank + i, party_size=PARTY_SIZE, kind="full") for i in range(n_rot)]
    pk = [ops.packed_to_device(r, dev) for r in recs]
    mask = torch.empty((B, A), dtype=torch.int32, device=dev)
    q = torch.empty((B, A), dtype=torch.float32, device=dev)
    act = torch.empty((B,), dtype=torch.int32, device=dev)

    train = args.mode == "train"
    if train:
        del ws
        ws = wide.train_workspace(nl, H, B, A, dev)
        g = torch.Generator().manual_seed(1)
        dq = (torch.randn((B, A), generator=g) / (B * A)).to(dev)  # d loss / d q of a mean-reduced loss, resident
        grad = torch.empty(N_params(nl, H, Cn), dtype=torch.float32, device=dev)

    def step(i):
        if train:
            wide.train_forward(pk[i % n_rot], A, prepared, nl, H, ws, out_q=q, out_mask=mask)
            wide.train_backward(prepared, nl, H, B, A, ws, dq, out_grad=grad)
        else:
            wide.wide_forward(pk[i % n_rot], A, prepared, nl, H, ws=ws, out_mask=mask, out_q=q, out_action=act)

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
        step(i)
    ev1.record()
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    barrier()
    dev_ms = max_over_ranks(ev0.elapsed_time(ev1))
    clocks = sampler.summary(t0, t1)
    checksum = float(grad.abs().sum().item()) if train else int(act.sum().item())

    if train:
        pin = torch.empty((B, 144), dtype=torch.uint8, pin_memory=True)
        dpk = torch.empty((B, 144), dtype=torch.uint8, device=dev)
        hq = torch.empty((B, A), dtype=torch.float32, pin_memory=True)

        def e2e_step(i):
            pin.numpy()[:] = np.ascontiguousarray(recs[i % n_rot]).view(np.uint8).reshape(B, 144)
            dpk.copy_(pin, non_blocking=True)
            wide.train_forward(dpk, A, prepared, nl, H, ws, out_q=q, out_mask=mask)
            wide.train_backward(prepared, nl, H, B, A, ws, dq, out_grad=grad)
            hq.copy_(q, non_blocking=True)  # the Q values the loss is formed from
            torch.cuda.current_stream(dev).synchronize()
        api = "host records -> pinned staging, H2D, pk_wide_train_forward, pk_wide_train_backward, D2H of Q, sync"
    else:
        runner = wide.WideRunner(model, A, dev)

        def e2e_step(i):
            runner.decide(recs[i % n_rot])
        api = "WideRunner.decide(host records): pinned staging, H2D, pk_wide_forward, D2H of Q and actions, sync"
    for i in range(3):
        e2e_step(i)
    torch.cuda.synchronize()
    barrier()
    te0 = time.perf_counter()
    for i in range(args.steps):
        e2e_step(i)
    te = max_over_ranks(time.perf_counter() - te0)
    barrier()
    sampler.stop()
    e2e = {"value": B * world * args.steps / te, "unit": UNIT, "h2d_bytes_per_step": B * 144,
           "d2h_bytes_per_step": B * A * 4 + (0 if train else B * 4), "ms_per_step": te / args.steps * 1e3, "api": api}
    cpu = None
    if rank == 0 and world == 1 and args.cpu_seconds > 0 and not train:
        cpu = cpu_baseline(nl, H, args.cpu_seconds)
    if rank == 0:
        ms = dev_ms / args.steps
        M = B * A
        flops = 2.0 * M * H * (Cn + (nl - 1) * H + 1)  # useful FLOPs: true K = 617 in layer 0, + the head
        if train:  # + dW of every layer (same FLOPs as its forward) + dX of every layer but the first + head
            flops += 2.0 * M * H * (Cn + (nl - 1) * H + 1) + 2.0 * M * H * (nl - 1) * H
        peak, src = tensor_peak()
        ach = flops / (ms * 1e-3) / 1e12
        line = {
            "metric": METRIC if not train else
            "samples/sec, wide net 4x2048 forward + backward (encode + tensor-core MLP + flat gradient)",
            "value": B * world * args.steps / (dev_ms * 1e-3), "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": ms, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "bf16 (fp32 accumulate)", "data": "synthetic",
            "config": {"workload": f"wide policy net {nl}x{H}, {B} synthetic battle states per GPU = {M} GEMM rows "
                                   f"(BASELINE configs[4]), " + ("forward + backward to the flat gradient (all columns)"
                                                                   if train else "forward + masked argmax"),
                       "states_per_gpu_per_step": B, "party_size": PARTY_SIZE, "action_slots": A, "channels": Cn,
                       "sharding": f"dp{world} (independent battles, no collective)",
                       "l2": f"activations are {M * H * 2 / 1e6:.0f} MB per layer (> the 126 MB L2); {n_rot} rotating inputs"},
            "roofline": {"bound": "tensor", "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak,
                         "traffic": None, "kernel": "pk_wide_gemm_kernel", "flops_per_step": flops, "peak_source": src},
            "e2e": e2e, "cpu_baseline": cpu, "clocks": clocks, "gpu_launches": args.steps * ((nl + 2) + (5 * nl + 4) if train else (nl + 3)), "mode": args.mode,
            "checksum": checksum,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
