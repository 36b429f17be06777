"""Secondary legs of bench.py: the other BASELINE.json configurations under the driver's clock.

bench.py keeps its headline (configs[1]: 4 096 states per launch, default MLP) exactly as the contract says and then
calls `run_all()`, which runs bounded legs (about a minute in total) whose results ride inside
`roofline.secondary` of the same JSON line:

  configs[2]  65 536 states sharded over the N GPUs (decision kernel, features written)
  configs[3]  DQN minibatch update, 32 768 transitions per step, STRONG scaling.  At N > 1 the gradient exchange --
              the path's only collective (trainer.py:187-195: mean loss, clamp after the reduce) -- is run by both
              routes (fused peer-memory sum + clamp + Adam kernel; NCCL allreduce + clamp + Adam) and checked first:
              three steps by each route from the same weights, parameters bit-identical on every rank (all-gather),
              routes within 1e-6 of each other and of the single-GPU full-batch update  -> `parity_ok`.
  configs[3]w the same update on the wide 4 x 2048 net (55 MB fp32 gradient, NCCL buckets overlapped with the backward)
  configs[4]  wide 4 x 2048 net, forward and forward + backward on the tensor cores; fractions of the burst AND the
              sustained measured bf16 peak
  configs[0]  64 concurrent battles through the lock-step driver (latency per decision round)
  dense       pk_mlp_forward on pre-encoded dense features (what agent.py:28-30 / party_evaluator_quick.py:101 hand in):
              65 536 states, HBM-bound by the 22 212 B + 36 B per decision it reads and writes
  gpu_reference  the reference's own GPU path (party_evaluator_quick.py:93-104): MLPModel in eager PyTorch on cuda on
              PRE-ENCODED dense features + mask + argmax, default and wide net -- the GPU bar SURVEY.md 2.2 names.

Every leg times with CUDA events on the launching stream after warm-up, max over ranks, and never raises into the
headline: a failing leg reports {"error": ...}.
"""
from __future__ import annotations

import json
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
A = 9
PARTY_SIZE = 6
ALG_BYTES_PER_DECISION = 144 + 617 * 9 * 4 + 36 + 36 + 4


def _peaks():
    d = {}
    f = ROOT / "MEASURED_PEAKS.json"
    if f.exists():
        try:
            d = json.loads(f.read_text())
        except Exception:
            d = {}
    return {"hbm": float(d.get("hbm_gbs", 6650.0)), "bf16_burst": float(d.get("bf16_tflops", 1637.2)),
            "bf16_sustained": float(d.get("bf16_tflops_sustained", 1434.9)),
            "source": "measured (MEASURED_PEAKS.json)" if d else "fallback (B200_PROFILING.md)"}


def _timed(ctx, fn, steps, warm=3):
    """CUDA events on the current stream around `steps` calls of fn(i); returns max-over-ranks ms per step."""
    import torch
    for i in range(warm):
        fn(i)
    torch.cuda.synchronize()
    ctx.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        fn(warm + i)
    e1.record()
    torch.cuda.synchronize()
    ctx.barrier()
    return ctx.max_over_ranks(e0.elapsed_time(e1)) / steps


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


# ---- configs[2]: 65 536 states over the N GPUs -------------------------------------------------------------------
def leg_shard65536(ctx, total=65536, steps=10):
    import torch
    from pokeai_b200 import ops
    from pokeai_b200.packed import synth_packed
    dev, world, rank = ctx.dev, ctx.world, ctx.rank
    B = total // world
    n_rot = 2 if B * 617 * A * 4 > 126 * 1024 * 1024 else 4
    pk = [ops.packed_to_device(synth_packed(B, seed=7000 + 10 * rank + i, party_size=PARTY_SIZE, kind="full"), dev)
          for i in range(n_rot)]
    feats = [torch.empty((B, 617, A), dtype=torch.float32, device=dev) for _ in range(n_rot)]
    mask = torch.empty((B, A), dtype=torch.int32, device=dev)
    q = torch.empty((B, A), dtype=torch.float32, device=dev)
    act = torch.empty((B,), dtype=torch.int32, device=dev)

    def step(i):
        ops.decide(pk[i % n_rot], A, ctx.prepared, 2, 64, out_feat=feats[i % n_rot], out_mask=mask, out_q=q, out_action=act)

    ms = _timed(ctx, step, steps)
    pk_ = _peaks()
    ach = ALG_BYTES_PER_DECISION * B / (ms * 1e-3) / 1e9
    del feats
    return {"config": f"configs[2] {total} states sharded over {world} GPU(s), {B} per launch per GPU, default MLP 2x64, "
                      "features+mask+Q+action written", "scaling": "strong", "decisions_per_s": B * world / (ms * 1e-3),
            "launch_us": ms * 1e3, "steps": steps, "roofline_bound": "hbm", "achieved_gbs_per_gpu": ach,
            "frac": ach / pk_["hbm"], "peak": pk_["hbm"], "l2": f"{n_rot} rotating output sets of {B * 617 * A * 4 / 1e6:.0f} MB"}


# ---- dense-input forward (Agent._calc_q_vector_batch's input) -------------------------------------------------------
def leg_dense_forward(ctx, total=65536, steps=10):
    import torch
    from pokeai_b200 import ops
    from pokeai_b200.packed import synth_packed
    dev, world, rank = ctx.dev, ctx.world, ctx.rank
    B = total // world
    n_rot = 2 if B * 617 * A * 4 > 126 * 1024 * 1024 else 4
    feats = []
    for i in range(n_rot):
        pk = ops.packed_to_device(synth_packed(B, seed=7100 + 10 * rank + i, party_size=PARTY_SIZE, kind="full"), dev)
        feats.append(ops.encode(pk, A)[0])
    q = torch.empty((B, A), dtype=torch.float32, device=dev)

    def step(i):
        ops.mlp_forward(ctx.prepared, 2, 64, feats[i % n_rot], out=q)

    ms = _timed(ctx, step, steps)
    pk_ = _peaks()
    bytes_per = 617 * A * 4 + A * 4
    ach = bytes_per * B / (ms * 1e-3) / 1e9
    del feats
    return {"config": f"dense-input forward (pk_mlp_forward, tensor cores): {total} states over {world} GPU(s), {B} per launch "
                      "per GPU, default MLP 2x64, dense (B, 617, 9) fp32 features in HBM -> Q", "scaling": "strong",
            "decisions_per_s": B * world / (ms * 1e-3), "launch_us": ms * 1e3, "steps": steps, "roofline_bound": "hbm",
            "bytes_per_decision": bytes_per, "achieved_gbs_per_gpu": ach, "frac": ach / pk_["hbm"], "peak": pk_["hbm"],
            "l2": f"{n_rot} rotating input sets of {B * 617 * A * 4 / 1e6:.0f} MB"}


# ---- configs[3]: the DQN update and its gradient exchange ---------------------------------------------------------
def leg_update(ctx, NG=32768, nl=2, H=64, steps=20):
    import numpy as np
    import torch
    from pokeai_b200 import _native as N, ops
    from pokeai_b200.dist import PeerGradExchange, shard_bounds
    from pokeai_b200.mlp_model import MLPModel
    dev, world, rank, dist = ctx.dev, ctx.world, ctx.rank, ctx.dist
    Cn = ops.n_channels()
    lo, hi = shard_bounds(NG, world, rank)
    n = hi - lo
    n_rot = 3
    full_sets, sets = [], []
    for i in range(n_rot):  # every rank generates the same global minibatch and keeps its contiguous shard
        arrs = synth_transitions(NG, 10 + i)

        def to_dev(a, b):
            s, s2, ac, r, t = (np.ascontiguousarray(x[a:b]) for x in arrs)
            return (ops.packed_to_device(s, dev), ops.packed_to_device(s2, dev), torch.from_numpy(ac).to(dev),
                    torch.from_numpy(r).to(dev), torch.from_numpy(t).to(dev))
        sets.append(to_dev(lo, hi))
        if world > 1:
            full_sets.append(to_dev(0, NG))
    torch.manual_seed(1)
    sd = MLPModel((Cn, A), A, n_layers=nl, n_channels=H).folded_state_dict()
    torch.manual_seed(2)
    sd_t = MLPModel((Cn, A), A, n_layers=nl, n_channels=H).folded_state_dict()
    prep_t = ops.prepare_weights(sd_t, dev)[0]
    lr, betas, eps, gamma = 1e-4, (0.9, 0.999), 1e-8, 0.95

    class Route:
        def __init__(self, name, nmax):
            self.name = name
            self.prep, _, _, self.flat = ops.prepare_weights(sd, dev)
            self.m, self.v = torch.zeros_like(self.flat), torch.zeros_like(self.flat)
            self.step_no = 0
            self.scratch = torch.empty(N.lib().pk_dqn_scratch_floats(nl, H, Cn, max(nmax, 1)), dtype=torch.float32, device=dev)
            self.loss = torch.zeros(1, dtype=torch.float32, device=dev)
            self.grad = torch.empty_like(self.flat)

    px = PeerGradExchange.create(sd_numel(sd), dev) if world > 1 else None

    def step(rt, batch, n_global):
        s, s2, ac, r, t = batch
        rt.step_no += 1
        if rt.name == "peer":
            gslot, lslot = px.next_slot()
            ops.dqn_grad(rt.prep, prep_t, nl, H, A, s, s2, ac, r, t, gamma, True, n_global=n_global, scratch=rt.scratch,
                         out_grad=gslot, out_loss=lslot)
            px.reduce_clamp_adam(rt.flat, rt.m, rt.v, nl, H, Cn, rt.step_no, lr, betas, eps, prepared=rt.prep)
            return
        ops.dqn_grad(rt.prep, prep_t, nl, H, A, s, s2, ac, r, t, gamma, True, n_global=n_global, scratch=rt.scratch,
                     out_grad=rt.grad, out_loss=rt.loss)
        if rt.name == "nccl":
            dist.all_reduce(rt.grad, op=dist.ReduceOp.SUM)
        ops.clamp_adam(rt.flat, rt.grad, rt.m, rt.v, nl, H, rt.step_no, lr=lr, betas=betas, eps=eps, prepared=rt.prep)

    # graph-replayable route: the whole step (control-block advance, next-state pass, backward, reduce, [peer exchange +]
    # clamp + Adam) captured once per gradient-buffer parity and replayed -- no launch arguments change between steps
    class GraphRoute(Route):
        def __init__(self, nmax):
            super().__init__("graph", nmax)
            self.ctl = ops.make_update_ctl(dev, lr, betas, eps)
            self.graphs = {}

        def capture(self, batches):
            """batches[p]: the (fixed) input set the graph of parity p reads"""
            side = torch.cuda.Stream(device=dev)
            for p in (1, 0):
                s, s2, ac, r, t = batches[p]

                def body():
                    ops.update_ctl_advance(self.ctl)
                    if px is not None:
                        gslot, lslot = px.slot(p)
                        ops.dqn_grad(self.prep, prep_t, nl, H, A, s, s2, ac, r, t, gamma, True, n_global=NG, scratch=self.scratch,
                                     out_grad=gslot, out_loss=lslot)
                        px.reduce_clamp_adam_ctl(p, self.flat, self.m, self.v, nl, H, Cn, self.ctl, prepared=self.prep)
                    else:
                        ops.dqn_grad(self.prep, prep_t, nl, H, A, s, s2, ac, r, t, gamma, True, n_global=NG, scratch=self.scratch,
                                     out_grad=self.grad, out_loss=self.loss)
                        ops.clamp_adam_ctl(self.flat, self.grad, self.m, self.v, nl, H, self.ctl, prepared=self.prep)
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=side):
                    body()
                self.graphs[p] = g

        def step(self):
            self.step_no += 1
            self.graphs[self.step_no & 1].replay()

    out = []
    base = {"config": f"configs[3] DQN minibatch update, {NG} packed transitions per step (global), default MLP {nl}x{H}, "
                      f"Double DQN, 5% terminal; dp{world}: contiguous shards of one global minibatch",
            "scaling": "strong", "unit": "transitions/s", "gradient_floats": sd_numel(sd)}
    if world == 1:
        rt = Route("single", n)
        ms = _timed(ctx, lambda i: step(rt, sets[i % n_rot], NG), steps)
        out.append(dict(base, exchange="none", transitions_per_s=NG / (ms * 1e-3), ms_per_step=ms, steps=steps))
        # the same step as a replayed CUDA graph, checked against three plain steps from the same weights first
        ref, gr = Route("single", n), GraphRoute(n)
        gr.capture({1: sets[0], 0: sets[1]})
        for k in range(3):
            step(ref, sets[k % 2], NG)
            gr.step()
        torch.cuda.synchronize()
        diff = float((ref.flat - gr.flat).abs().max())
        ms = _timed(ctx, lambda i: gr.step(), steps)
        out.append(dict(base, exchange="none; the step replayed as one CUDA graph (device-side Adam step count)",
                        transitions_per_s=NG / (ms * 1e-3), ms_per_step=ms, steps=steps, parity_ok=bool(diff <= 1e-7),
                        parity={"graph_vs_plain_max_abs": diff}))
        return out
    # ---- parity first: both routes and the single-GPU full-batch update, three steps from the same weights -------
    routes = {"nccl": Route("nccl", n), "full": Route("full", NG)}
    if px is not None:
        routes["peer"] = Route("peer", n)
    for k in range(3):
        step(routes["nccl"], sets[k % n_rot], NG)
        step(routes["full"], full_sets[k % n_rot], NG)
        if px is not None:
            step(routes["peer"], sets[k % n_rot], NG)
    torch.cuda.synchronize()
    checks = {}
    for name in ("nccl", "peer"):
        if name not in routes:
            continue
        flat = routes[name].flat
        allp = [torch.empty_like(flat) for _ in range(world)]
        dist.all_gather(allp, flat)
        checks[f"{name}_bit_identical_across_ranks"] = bool(all(torch.equal(allp[0], p) for p in allp))
        checks[f"{name}_vs_single_gpu_max_abs"] = float((flat - routes["full"].flat).abs().max())
    if px is not None:
        checks["peer_vs_nccl_max_abs"] = float((routes["peer"].flat - routes["nccl"].flat).abs().max())
        checks["peer_timeout_flag"] = float(px.loss_status[1])
    ok = all(v for k, v in checks.items() if k.endswith("across_ranks")) and \
        all(v <= 1e-6 for k, v in checks.items() if k.endswith("max_abs")) and checks.get("peer_timeout_flag", 0.0) == 0.0
    okt = torch.tensor([1 if ok else 0], dtype=torch.int32, device=dev)
    dist.all_reduce(okt, op=dist.ReduceOp.MIN)
    parity_ok = bool(int(okt.item()) == 1)
    del full_sets
    for name in ("peer", "nccl"):
        if name not in routes:
            out.append(dict(base, exchange=name, error="peer memory not mappable on this box"))
            continue
        rt = routes[name]
        ms = _timed(ctx, lambda i: step(rt, sets[i % n_rot], NG), steps)
        out.append(dict(base, exchange=("peer: gradient sum over NVLink peer memory + clamp + Adam in one kernel "
                                        "(pk_peer_clamp_adam)" if name == "peer" else
                                        "nccl: one sum-allreduce of the flat gradient, then pk_clamp_adam"),
                        transitions_per_s=NG / (ms * 1e-3), ms_per_step=ms, steps=steps, parity_ok=parity_ok,
                        parity=checks, ranks=world))
    if px is not None:
        # the peer route replayed as CUDA graphs (one per buffer parity): epoch and Adam step advance on the device.
        # Continues the peer exchange's epochs: the control block starts at the epoch the plain route has reached.
        peer = routes["peer"]
        gr = GraphRoute(n)
        gr.flat.copy_(peer.flat)
        gr.m.copy_(peer.m)
        gr.v.copy_(peer.v)
        gr.prep.copy_(peer.prep)
        gr.ctl = ops.make_update_ctl(dev, lr, betas, eps, step=peer.step_no, epoch=px.epoch)
        gr.step_no = px.epoch  # parity of the next epoch selects the graph
        ref = Route("nccl", n)
        for t_, s_ in ((ref.flat, peer.flat), (ref.m, peer.m), (ref.v, peer.v), (ref.prep, peer.prep)):
            t_.copy_(s_)
        ref.step_no = peer.step_no
        nxt = (px.epoch + 1) & 1  # parity of the first graph step
        gr.capture({nxt: sets[0], nxt ^ 1: sets[1]})
        for k in range(4):
            step(ref, sets[k % 2], NG)
            gr.step()
        px.epoch += 4
        torch.cuda.synchronize()
        allp = [torch.empty_like(gr.flat) for _ in range(world)]
        dist.all_gather(allp, gr.flat)
        gchecks = {"graph_bit_identical_across_ranks": bool(all(torch.equal(allp[0], p) for p in allp)),
                   "graph_vs_nccl_max_abs": float((gr.flat - ref.flat).abs().max()),
                   "peer_timeout_flag": float(px.loss_status[1])}
        gok = gchecks["graph_bit_identical_across_ranks"] and gchecks["graph_vs_nccl_max_abs"] <= 1e-6 and \
            gchecks["peer_timeout_flag"] == 0.0
        okt = torch.tensor([1 if gok else 0], dtype=torch.int32, device=dev)
        dist.all_reduce(okt, op=dist.ReduceOp.MIN)
        ms = _timed(ctx, lambda i: gr.step(), steps + (steps & 1), warm=4)  # whole graph pairs
        px.epoch += steps + (steps & 1) + 4
        out.append(dict(base, exchange="peer, the step replayed as CUDA graphs (one per buffer parity; epoch and Adam step "
                                       "advance on the device: pk_update_ctl)",
                        transitions_per_s=NG / (ms * 1e-3), ms_per_step=ms, steps=steps + (steps & 1),
                        parity_ok=bool(int(okt.item()) == 1), parity=gchecks, ranks=world))
    return out


# ---- the same update with bn: true (train-mode BatchNorm, csrc/pk_bn.cuh): single GPU, plain FP32 kernels ----------
def leg_bn_update(ctx, NG=4096, nl=2, H=64, steps=10):
    """Trainer._update_from_arrays of a bn=True network through the public host API (host arrays in, loss out, so every
    step includes its host->device copies): a reported figure for a path built for coverage, not tuned."""
    if ctx.world > 1:
        return None  # batch statistics are not synchronised across ranks: Trainer refuses a data-parallel bn update
    import numpy as np
    import torch
    from pokeai_b200.trainer import Trainer
    torch.manual_seed(11)
    tr = Trainer({"n_layers": nl, "n_channels": H, "bn": True}, {"batch_size": NG, "lr": 1e-4},
                 {"party_size": PARTY_SIZE}, device=ctx.dev)
    sets = [synth_transitions(NG, 40 + i) for i in range(2)]
    losses = []

    def step(i):
        s, s2, ac, r, t = sets[i % 2]
        losses.append(tr._update_from_arrays(s, s2, ac, r, t))

    ms = _timed(ctx, step, steps)
    return {"config": f"DQN update with bn: true (train-mode BatchNorm in all three forwards, mlp_model.py:12-15): {NG} "
                      f"transitions, MLP {nl}x{H}, Double DQN, through Trainer._update_from_arrays (host arrays in, loss out)",
            "transitions_per_s": NG / (ms * 1e-3), "update_ms": ms, "steps": steps, "scaling": "single GPU only",
            "loss_finite": bool(np.all(np.isfinite(losses))),
            "parity": "tests/test_gpu_bn.py: golden vectors of the reference Trainer._update with bn: true; torch fp64"}


def sd_numel(sd):
    return int(sum(v.numel() for v in sd.values()))


# ---- configs[3] on the wide net: 55 MB gradient exchange -----------------------------------------------------------
def leg_wide_update(ctx, NG=32768, nl=4, H=2048, steps=4):
    import numpy as np
    import torch
    from pokeai_b200 import ops
    from pokeai_b200.dist import shard_bounds
    from pokeai_b200.trainer import Trainer
    dev, world, rank, dist = ctx.dev, ctx.world, ctx.rank, ctx.dist
    lo, hi = shard_bounds(NG, world, rank)
    torch.manual_seed(0)
    tr = Trainer({"n_layers": nl, "n_channels": H, "bn": False}, {"batch_size": NG, "lr": 1e-4},
                 {"party_size": PARTY_SIZE}, device=dev)
    sets = []
    for i in range(2):
        s, s2, ac, r, t = (np.ascontiguousarray(x[lo:hi]) for x in synth_transitions(NG, 30 + i))
        sets.append((ops.packed_to_device(s, dev), ops.packed_to_device(s2, dev), torch.from_numpy(ac).to(dev),
                     torch.from_numpy(r).to(dev), torch.from_numpy(t).to(dev)))

    fwd = H * (617 + (nl - 1) * H + 1)
    fma = 2 * 9 * fwd + fwd + fwd + (nl - 1) * H * H  # two 9-column forwards, 1-column forward, dW, dX
    pk_ = _peaks()
    n_par = tr._flat.numel()

    def measure(route):
        px = tr._peer_exchange_wide(dist) if (route == "peer" and world > 1) else None
        if route == "peer" and world > 1 and px is None:
            return None

        def step(i):
            if px is not None:  # shard gradient straight into the exchange buffer, ONE kernel sums it over the ranks
                grad, loss_slot = px.next_slot()
                loss_slot.copy_(tr._wide_grad_device(*sets[i % 2], NG, None, out_grad=grad))
                px.allreduce()
                tr._wide_apply(grad)
            else:
                tr._wide_grad_device(*sets[i % 2], NG, dist if world > 1 else None)  # per-layer NCCL buckets inside at N > 1
                tr._wide_apply()

        ms = _timed(ctx, step, steps, warm=2)
        checks = None
        if world > 1:
            allp = [torch.empty_like(tr._flat) for _ in range(world)]
            dist.all_gather(allp, tr._flat)
            checks = {"parameters_bit_identical_across_ranks": bool(all(torch.equal(allp[0], p) for p in allp))}
            if px is not None:
                checks["peer_timeout_flag"] = float(px.loss_status[1].item())
        tf = 2.0 * fma * (hi - lo) / (ms * 1e-3) / 1e12
        exch = ("none" if world == 1 else
                "peer: sliced reduce-scatter + all-gather over NVLink peer memory in one kernel (pk_peer_allreduce), then "
                "clamp + Adam" if px is not None else "nccl: one allreduce bucket per layer started from inside the backward")
        return {"config": f"configs[3] on the wide net {nl}x{H}: {NG} transitions per step (global), dp{world}",
                "scaling": "strong", "unit": "transitions/s", "transitions_per_s": NG / (ms * 1e-3), "ms_per_step": ms,
                "steps": steps, "gradient_floats": n_par, "exchange": exch, "parity": checks,
                "parity_ok": None if checks is None else bool(checks["parameters_bit_identical_across_ranks"]
                                                              and checks.get("peer_timeout_flag", 0.0) == 0.0),
                "tflops_per_gpu": tf, "frac_burst": tf / pk_["bf16_burst"], "frac_sustained": tf / pk_["bf16_sustained"]}

    out = []
    for route in (("peer", "nccl") if world > 1 else ("nccl",)):
        r = measure(route)
        if r is not None:
            out.append(r)
    del tr
    return out


# ---- configs[4]: wide net on the tensor cores --------------------------------------------------------------------
def leg_wide(ctx, B=4096, nl=4, H=2048, steps=10):
    import torch
    from pokeai_b200 import ops, wide
    from pokeai_b200.mlp_model import MLPModel
    from pokeai_b200.packed import synth_packed
    dev, world, rank = ctx.dev, ctx.world, ctx.rank
    Cn = ops.n_channels()
    torch.manual_seed(0)
    model = MLPModel((Cn, A), A, n_layers=nl, n_channels=H).eval()
    prepared, _, _ = wide.prepare_wide(model.folded_state_dict(), dev)
    pk = [ops.packed_to_device(synth_packed(B, seed=2000 + 10 * rank + i, party_size=PARTY_SIZE, kind="full"), dev)
          for i in range(3)]
    mask = torch.empty((B, A), dtype=torch.int32, device=dev)
    q = torch.empty((B, A), dtype=torch.float32, device=dev)
    act = torch.empty((B,), dtype=torch.int32, device=dev)
    ws = wide.workspace(nl, H, B, A, dev)
    ms_f = _timed(ctx, lambda i: wide.wide_forward(pk[i % 3], A, prepared, nl, H, ws=ws, out_mask=mask, out_q=q,
                                                   out_action=act), steps)
    del ws
    wt = wide.train_workspace(nl, H, B, A, dev)
    g = torch.Generator().manual_seed(1)
    dq = (torch.randn((B, A), generator=g) / (B * A)).to(dev)
    grad = torch.empty(H * Cn + H + (nl - 1) * (H * H + H) + H + 1, dtype=torch.float32, device=dev)

    def train(i):
        wide.train_forward(pk[i % 3], A, prepared, nl, H, wt, out_q=q, out_mask=mask)
        wide.train_backward(prepared, nl, H, B, A, wt, dq, out_grad=grad)

    ms_t = _timed(ctx, train, steps)
    M = B * A
    f_fwd = 2.0 * M * H * (Cn + (nl - 1) * H + 1)
    f_trn = 2.0 * f_fwd + 2.0 * M * H * (nl - 1) * H
    pk_ = _peaks()
    out = []
    for name, ms, fl, unit in (("forward + masked argmax", ms_f, f_fwd, "decisions/s"),
                               ("forward + backward to the flat gradient (all columns)", ms_t, f_trn, "samples/s")):
        tf = fl / (ms * 1e-3) / 1e12
        out.append({"config": f"configs[4] wide net {nl}x{H}, {B} states per GPU = {M} GEMM rows, {name}; dp{world}, no collective",
                    "scaling": "weak", "unit": unit, "per_s": B * world / (ms * 1e-3), "ms_per_step": ms, "steps": steps,
                    "roofline_bound": "tensor", "tflops_per_gpu": tf, "frac_burst": tf / pk_["bf16_burst"],
                    "frac_sustained": tf / pk_["bf16_sustained"], "peak_burst": pk_["bf16_burst"],
                    "peak_sustained": pk_["bf16_sustained"], "dtype": "bf16 operands, fp32 accumulate (tcgen05, TMEM)"})
    return out


# ---- the reference's own GPU path: eager PyTorch on pre-encoded features -------------------------------------------
def leg_gpu_reference(ctx, B=4096, steps=10):
    """party_evaluator_quick.py:93-104: model(torch.from_numpy(batch).to(device)) -> .to('cpu') -> mask -> -inf -> max /
    argmax.  Timed twice: device-resident (features already in HBM: the most favourable reading) and as the reference
    calls it (dense features host -> device every call, Q back to the host)."""
    import numpy as np
    import torch
    from pokeai_b200 import ops
    from pokeai_b200.packed import synth_packed
    if ctx.rank != 0:
        return None
    dev = ctx.dev
    kind, RefMLP = "port", None
    try:
        from oracle import ref_stage
        R = ref_stage.load()
        if R is not None:
            RefMLP, kind = R.MLPModel, "reference"
    except Exception:
        RefMLP = None
    if RefMLP is None:  # the same module stack written with torch.nn (mlp_model.py:5-30)
        class RefMLP(torch.nn.Module):
            def __init__(self, input_shape, output_dim, n_layers=2, n_channels=64, bn=False):
                super().__init__()
                cur, layers = input_shape[0], []
                for _ in range(n_layers):
                    layers.append(torch.nn.Conv1d(cur, n_channels, 1))
                    cur = n_channels
                self.layers = torch.nn.ModuleList(layers)
                self.output = torch.nn.Conv1d(cur, 1, 1)

            def forward(self, x):
                h = x
                for layer in self.layers:
                    h = torch.relu(layer(h))
                return self.output(h).view(h.shape[0], -1)
    pk = ops.packed_to_device(synth_packed(B, seed=0, party_size=PARTY_SIZE, kind="full"), dev)
    feat, mask = ops.encode(pk, A)  # pre-encoded features: the reference's encoder is CPU Python and is NOT timed here
    torch.cuda.synchronize()
    feat_host = feat.cpu().numpy()
    mask_host = mask.cpu().numpy()
    out = []
    for nl, H in ((2, 64), (4, 2048)):
        torch.manual_seed(0)
        model = RefMLP((617, A), A, n_layers=nl, n_channels=H).to(dev).eval()

        def resident(i):
            with torch.no_grad():
                qv = model(feat)
                qv[mask == 0] = -float("inf")
                return torch.argmax(qv, dim=1)

        def as_called(i):
            with torch.no_grad():
                qv = model(torch.from_numpy(feat_host).to(dev)).to("cpu").numpy()
            qv[mask_host == 0] = -np.inf
            return np.argmax(qv, axis=1)

        ms_r = _timed(ctx_single(ctx), resident, steps)
        t0 = time.perf_counter()
        n_call = 3
        as_called(0)
        t0 = time.perf_counter()
        for i in range(n_call):
            as_called(i)
        ms_c = (time.perf_counter() - t0) / n_call * 1e3
        out.append({"config": f"gpu_reference: the reference's MLPModel {nl}x{H} in eager PyTorch on cuda:0 on PRE-ENCODED "
                              f"dense features ({B} states) + mask + argmax (party_evaluator_quick.py:93-104); fp32, torch "
                              f"defaults (cudnn.allow_tf32={torch.backends.cudnn.allow_tf32})",
                    "kind": kind, "unit": "decisions/s", "device_resident_per_s": B / (ms_r * 1e-3),
                    "device_resident_ms": ms_r, "as_called_per_s": B / (ms_c * 1e-3), "as_called_ms": ms_c,
                    "as_called_h2d_bytes": int(feat_host.nbytes), "encode": "not included (CPU Python in the reference)"})
        del model
    return out


def ctx_single(ctx):
    """timing context for a rank-0-only leg: no collective inside"""
    from types import SimpleNamespace
    return SimpleNamespace(barrier=lambda: None, max_over_ranks=lambda x: x)


# ---- configs[0]: 64 concurrent battles through the lock-step driver ------------------------------------------------
def leg_concurrent64(ctx, n_battles=64, rounds=200):
    import numpy as np
    if ctx.rank != 0:
        return None
    from pokeai_b200.lockstep import LockStepDriver, SyntheticBattle
    from pokeai_b200.agent import AgentVal
    agent = AgentVal(ctx.model, ctx.fe)
    battles = [SyntheticBattle(seed=100 + i, party_size=PARTY_SIZE) for i in range(n_battles)]
    warm = 20
    for b in battles:
        b.pregenerate(rounds + warm)  # the stand-in for simulator + protocol parser is not part of the path: not timed
    sample = [b._queue[warm] for b in battles]  # one round's requests, for the CPU figure below
    drv = LockStepDriver(agent, battles)
    for _ in range(warm):
        drv.step()
    drv.time_pack = drv.time_gpu = 0.0
    drv.rounds = 0
    t0 = time.perf_counter()
    n_dec = 0
    for _ in range(rounds):
        n_dec += drv.step()
    dt = time.perf_counter() - t0
    t_pack, t_gpu = drv.time_pack / max(1, drv.rounds), drv.time_gpu / max(1, drv.rounds)
    drv.close()
    out = {"config": f"configs[0] {n_battles} concurrent battles in lock step (LockStepDriver: every pending request packed "
                     "straight into pinned records, ONE pk_decide launch per round reading them over PCIe, keys handed back "
                     "in order), default MLP, synthetic requests drawn beforehand",
           "unit": "decisions/s", "decisions_per_s": n_dec / dt, "round_trip_us": dt / rounds * 1e6,
           "us_per_decision": dt / max(1, n_dec) * 1e6, "rounds": rounds, "host_pack_us_per_round": t_pack * 1e6,
           "gpu_call_us_per_round": t_gpu * 1e6,
           "note": "latency-bound: 64 states fill 64 of 2 368 warp slots; the GPU call is launch + completion latency, "
                   "the rest of the round is the host side (request -> record: "
                   + ("the C packer csrc/pkpack.c, ~5 us" if _c_packer_in_use() else "Python struct packer, ~13 us")
                   + " per request, plus the policy callbacks in Python)"}
    # the reference serves the same 64 battles one at a time (sim.py:54-99): get_possible_actions, RLPolicyObservation,
    # FeatureExtractor.transform, MLPModel at batch 1, argmax -- timed here with its own classes when they are staged
    try:
        import torch
        from oracle import ref_port, ref_stage
        R = ref_stage.load()
        if R is not None:
            from pokeai.ai.common import get_possible_actions
            from pokeai.ai.generic_move_model.agent_val import AgentVal as RefAgentVal
            fe = R.FeatureExtractor(PARTY_SIZE)
            sd = {k: v.detach().cpu() for k, v in ctx.model.state_dict().items()}
            model = R.MLPModel(fe.input_shape, fe.output_dim, n_layers=ctx.model.n_layers, n_channels=ctx.model.n_channels,
                               bn=False)
            model.load_state_dict(sd)
            model.eval()
            ragent = RefAgentVal(model, fe)
            torch.set_num_threads(1)  # the reference's own setting for batch-1 inference (experiment_202009.md:48-58)

            def one_round():
                for bs, request in sample:
                    pas = get_possible_actions(bs, request)
                    if len(pas) > 1:
                        ragent.act(R.RLPolicyObservation(bs, request, pas), 0.0)
            with torch.no_grad():
                one_round()
                t0 = time.perf_counter()
                for _ in range(5):
                    one_round()
                out["cpu_reference_round_us"] = (time.perf_counter() - t0) / 5 * 1e6
            out["cpu_reference"] = "the reference's RLPolicy path (oracle/_ref), 64 decisions one at a time, 1 thread"
    except Exception as e:  # a baseline figure never takes the leg down
        out["cpu_reference_error"] = f"{type(e).__name__}: {e}"[:200]
    return out


def _c_packer_in_use():
    from pokeai_b200 import protocol
    return protocol._C_PACK is not None


def run_all(ctx, skip=()):
    legs = [("shard65536", leg_shard65536), ("dense", leg_dense_forward), ("update", leg_update), ("bn_update", leg_bn_update), ("wide", leg_wide), ("wide_update", leg_wide_update),
            ("gpu_reference", leg_gpu_reference), ("concurrent64", leg_concurrent64)]
    import torch
    out = []
    for name, fn in legs:
        if name in skip:
            continue
        t0 = time.perf_counter()
        try:
            res = fn(ctx)
        except Exception as e:  # a secondary leg never takes the headline down
            res = {"config": name, "error": f"{type(e).__name__}: {e}"[:400]}
        torch.cuda.synchronize()
        torch.cuda.empty_cache()
        if res is None:
            continue
        for r in (res if isinstance(res, list) else [res]):
            r["leg"] = name
            r["leg_wall_s"] = round(time.perf_counter() - t0, 2)
            out.append(r)
    return out
