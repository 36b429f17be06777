#!/usr/bin/env python
"""bench.py -- policy decisions/sec of the decision hot path (encode + MLP + masked select).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W

One "step" = one pass of the hot path over one batch of synthetic battle states:
BASELINE.json configs[1], 4096 Gen-1 6v6 states (A = 9 action slots, C = 617 channels),
default MLP 2 x 64, features and masks materialised in HBM, Q values and the masked-argmax
action written.  At N > 1 every rank runs the same per-GPU batch on its own shard of
states (battles are independent; no data-path collective) => weak scaling.

Prints ONE JSON line (rank 0).  Keys beyond the base contract:
  roofline      achieved HBM GB/s of the decision kernel = algorithmic bytes per launch
                (22 432 B per decision) / CUDA-event time per launch.
  e2e           the same metric through the host-buffer public API, every step moving its records host ->
                device and its Q + actions device -> host: `value` = DecisionPipeline (three steps in flight,
                records in pinned host memory read by the kernel over PCIe); `pageable_input_value` = the same
                from ordinary host memory (one more host copy per step); `sync_value` = one blocking
                AgentVal.act_batch call per step.
  cpu_baseline  the reference algorithm's CPU path (oracle port of the reference's Python,
                oracle/ref_port.py) timed on this box's host cores on the same states.
  clocks        SM clock / throttle reasons sampled through NVML during the timed region.

--impl reference times that CPU path alone (rank 0 only) and prints the same line shape.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

METRIC = "policy decisions/sec (encode+MLP+masked sample)"
UNIT = "decisions/s"
PARTY_SIZE = 6
A = 9
ALG_BYTES_PER_DECISION = 144 + 617 * 9 * 4 + 36 + 36 + 4  # = 22 432
L2_BYTES = 126 * 1024 * 1024


def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=400)
    p.add_argument("--warmup", type=int, default=20)
    p.add_argument("--impl", default="b200", choices=["b200", "reference"])
    p.add_argument("--batch", type=int, default=4096, help="battle states per GPU per step")
    p.add_argument("--layers", type=int, default=2)
    p.add_argument("--channels", type=int, default=64)
    p.add_argument("--cpu-seconds", type=float, default=12.0, help="CPU baseline budget (0 = skip)")
    p.add_argument("--no-e2e", action="store_true")
    p.add_argument("--no-graph", action="store_true",
                   help="issue every launch from Python instead of replaying a CUDA graph of launches")
    return p.parse_args()


def peaks():
    f = ROOT / "MEASURED_PEAKS.json"
    if f.exists():
        try:
            return float(json.loads(f.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def workload_config(args, world):
    return {
        "workload": f"Gen-1 6v6 batched inference, {args.batch} synthetic battle states per GPU, "
                    f"default MLP {args.layers}x{args.channels}, features+mask+Q+action written (BASELINE configs[1])",
        "states_per_gpu_per_step": args.batch, "global_states_per_step": args.batch * world,
        "party_size": PARTY_SIZE, "action_slots": A, "channels": 617,
        "mlp": f"{args.layers}x{args.channels} fp32", "sharding": f"dp{world} (independent battles, no collective)",
    }


# ---- NVML clock sampler (runs during the timed region) --------------------------------------------
class ClockSampler:
    """Samples SM clock and throttle reasons of one GPU through NVML every ~2 ms from a thread.
    The recipe's `nvidia-smi -lms 200` is too coarse for a timed region of tens of milliseconds;
    NVML is what nvidia-smi reads."""

    BAD = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown")

    def __init__(self, index: int):
        self.samples = []  # (t, sm_mhz, reasons_bitmask)
        self.sm_max = None
        self._stop = threading.Event()
        self._thr = None
        self._marks = []
        try:
            import pynvml
            self.nv = pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            if vis:
                ids = [v.strip() for v in vis.split(",") if v.strip()]
                tok = ids[index] if index < len(ids) else str(index)
                self.h = (pynvml.nvmlDeviceGetHandleByIndex(int(tok)) if tok.isdigit()
                          else pynvml.nvmlDeviceGetHandleByUUID(tok))
            else:
                self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.sm_max = int(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception as e:  # no NVML: report that, never fake a clock
            self.nv = None
            self.err = repr(e)

    def _read(self):
        nv = self.nv
        mhz = int(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
        try:
            r = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
        except Exception:
            r = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
        return mhz, r

    def _run(self):
        while not self._stop.is_set():
            try:
                mhz, r = self._read()
                self.samples.append((time.perf_counter(), mhz, r))
            except Exception:
                pass
            time.sleep(0.002)

    def start(self):
        if self.nv is not None:
            self._thr = threading.Thread(target=self._run, daemon=True)
            self._thr.start()

    def stop(self):
        self._stop.set()
        if self._thr is not None:
            self._thr.join()

    def reasons(self, mask: int):
        nv = self.nv
        names = {
            "gpu_idle": getattr(nv, "nvmlClocksEventReasonGpuIdle", 0x1),
            "applications_clocks_setting": getattr(nv, "nvmlClocksEventReasonApplicationsClocksSetting", 0x2),
            "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4),
            "hw_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8),
            "sync_boost": getattr(nv, "nvmlClocksEventReasonSyncBoost", 0x10),
            "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
            "hw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40),
            "hw_power_brake_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwPowerBrakeSlowdown", 0x80),
            "display_clock_setting": getattr(nv, "nvmlClocksEventReasonDisplayClockSetting", 0x100),
        }
        return sorted(k for k, bit in names.items() if mask & bit)

    def summary(self, t0: float, t1: float):
        if self.nv is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "error": self.err}
        inside = [s for s in self.samples if t0 <= s[0] <= t1]
        note = "sampled inside the timed region"
        if not inside:  # region shorter than the sampling period: nearest samples around it
            inside = sorted(self.samples, key=lambda s: min(abs(s[0] - t0), abs(s[0] - t1)))[:3]
            note = "timed region shorter than the 2 ms sampling period; nearest samples"
        mhz = sorted(s[1] for s in inside)
        mask = 0
        for s in inside:
            mask |= s[2]
        reasons = [r for r in self.reasons(mask) if r != "gpu_idle"]
        return {"sm_mhz": mhz[len(mhz) // 2] if mhz else None, "sm_max_mhz": self.sm_max, "reasons": reasons,
                "samples": len(inside), "rejected": any(r in self.BAD for r in reasons), "note": note}


# ---- the reference algorithm on the host cores (oracle port; also the --impl reference arm) ----------
def cpu_decide_factory(rec, n_layers, n_channels, threads):
    """Returns (step(n) -> decisions done, description).  This is the ONLY place bench.py touches
    oracle/: the timed CPU baseline.  It runs the reference's algorithm the way the reference runs it:
    a Python loop of FeatureExtractor.transform over observation objects, np.stack, one batched
    MLPModel forward (torch CPU fp32, all host threads), mask -> -inf, argmax
    (feature_extractor.py:40-52, agent.py:28-35)."""
    import numpy as np
    import torch
    from oracle import ref_port
    from pokeai_b200.packed import unpack_observation
    from pokeai_b200.tables import load_tables
    T = load_tables()
    torch.set_num_threads(threads)
    obs = [unpack_observation(r, T) for r in rec]  # the reference receives objects; not timed
    sd = ref_port.init_mlp_params(T.n_channels, n_layers, n_channels, seed=0)

    def step(n):
        feats, masks = ref_port.transform_batch(obs[:n], PARTY_SIZE, T)
        with torch.no_grad():
            q = ref_port.mlp_forward(sd, torch.from_numpy(feats)).numpy()
        act = ref_port.masked_argmax(q, masks)
        return int(act.shape[0])

    return step


def host_threads():
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except Exception:
        return os.cpu_count() or 1


def run_cpu_baseline(rec, args, budget_s):
    threads = host_threads()
    step = cpu_decide_factory(rec, args.layers, args.channels, threads)
    n = min(len(rec), 256)
    t0 = time.perf_counter()
    step(n)  # warm-up on a slice, also an estimate of the rate
    rate = n / (time.perf_counter() - t0)
    n = int(min(len(rec), max(256, rate * budget_s / 3)))
    times = []
    t_all = time.perf_counter()
    while len(times) < 3 or (time.perf_counter() - t_all < budget_s and len(times) < 50):
        t0 = time.perf_counter()
        step(n)
        times.append(time.perf_counter() - t0)
    times.sort()
    med = times[len(times) // 2]
    return {"value": n / med, "unit": UNIT, "cores": threads, "kind": "port",
            "sample": f"{n} of the {len(rec)} states of one step, {len(times)} passes, median; encode loop is "
                      f"single-threaded Python as in the reference, torch forward uses {threads} threads"}


def reference_arm(args, rank, world):
    """--impl reference: the reference's CPU implementation of the path on the host cores.
    The reference is pure Python (nothing to compile into oracle/_ref) and /root/reference does not
    exist on the GPU box, so this is the oracle port, pinned to the reference by tests/golden/."""
    if rank != 0:
        return
    from pokeai_b200.packed import synth_packed
    rec = synth_packed(args.batch, seed=0, party_size=PARTY_SIZE, kind="full")
    threads = host_threads()
    step = cpu_decide_factory(rec, args.layers, args.channels, threads)
    # size the per-step sample so that (warmup + steps) passes end within ~150 s
    n = min(len(rec), 256)
    t0 = time.perf_counter()
    step(n)
    rate = n / (time.perf_counter() - t0)
    per_step = int(min(len(rec), max(64, rate * 150.0 / max(1, args.steps + args.warmup))))
    for _ in range(args.warmup):
        step(per_step)
    t0 = time.perf_counter()
    done = 0
    for _ in range(args.steps):
        done += step(per_step)
    dt = time.perf_counter() - t0
    value = done / dt
    sample = (f"each step = the first {per_step} of the {len(rec)} states of the workload; Python encode loop "
              f"(single-threaded, as the reference) + torch CPU fp32 forward on {threads} threads + masked argmax")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, max(1, args.gpus)),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "CPU arm runs on rank 0's host cores only; it does not scale with --gpus",
    }
    print(json.dumps(line), flush=True)


# ---- the B200 arm -----------------------------------------------------------------------------------
def b200_arm(args, rank, local_rank, world):
    import numpy as np
    import torch
    import torch.distributed as dist
    from pokeai_b200 import ops
    from pokeai_b200.agent import AgentVal
    from pokeai_b200.feature_extractor import FeatureExtractor
    from pokeai_b200.mlp_model import MLPModel
    from pokeai_b200.packed import synth_packed

    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl b200 needs a CUDA device (pokeai_b200 has no CPU path)")
    torch.cuda.set_device(local_rank)
    dev = torch.device(f"cuda:{local_rank}")
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    B, nl, H = args.batch, args.layers, args.channels
    fe = FeatureExtractor(PARTY_SIZE, device=dev)
    Cn = fe.input_shape[0]
    torch.manual_seed(0)
    model = MLPModel(fe.input_shape, fe.output_dim, n_layers=nl, n_channels=H).eval()  # random init, as torch does
    prepared = model.prepared(dev)

    # rotating buffers: the outputs of consecutive steps cover more than 2x the L2, so no step finds its
    # lines resident; every rank has its own shard of states (seeded by rank).
    feat_bytes = B * Cn * A * 4
    n_rot = max(3, int(np.ceil(2.0 * L2_BYTES / feat_bytes)) + 1)
    recs = [synth_packed(B, seed=1000 * rank + i, party_size=PARTY_SIZE, kind="full") for i in range(n_rot)]
    pk = [ops.packed_to_device(r, dev) for r in recs]
    feats = [torch.empty((B, Cn, A), dtype=torch.float32, device=dev) for _ in range(n_rot)]
    masks = [torch.empty((B, A), dtype=torch.int32, device=dev) for _ in range(n_rot)]
    qs = [torch.empty((B, A), dtype=torch.float32, device=dev) for _ in range(n_rot)]
    acts = [torch.empty((B,), dtype=torch.int32, device=dev) for _ in range(n_rot)]

    def step(i):
        j = i % n_rot
        ops.decide(pk[j], A, prepared, nl, H, out_feat=feats[j], out_mask=masks[j], out_q=qs[j], out_action=acts[j])

    # The timed steps are replayed from a CUDA graph holding G consecutive launches (the rotation over the output
    # sets included): same kernels, same buffers, same count, but no host work between launches, so the number
    # is the GPU's and not the Python launch loop's (which the NVML sampler thread slows further through the GIL).
    side = torch.cuda.Stream(device=dev)
    graph, G = None, 4 * n_rot
    sampler = ClockSampler(local_rank)
    sampler.start()
    with torch.cuda.stream(side):
        for i in range(max(3, args.warmup)):
            step(i)
        side.synchronize()
        if not args.no_graph and args.steps >= G:
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=side):
                for i in range(G):
                    step(i)
            graph.replay()  # one untimed replay
            side.synchronize()
        barrier()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        side.synchronize()
        t0 = time.perf_counter()
        ev0.record(side)
        done = 0
        if graph is not None:
            for _ in range(args.steps // G):
                graph.replay()
            done = (args.steps // G) * G
        for i in range(done, args.steps):  # remainder (or everything with --no-graph): plain launches
            step(i)
        ev1.record(side)
        side.synchronize()
        t1 = time.perf_counter()
    barrier()
    dev_ms = max_over_ranks(ev0.elapsed_time(ev1))
    clocks = sampler.summary(t0, t1)
    checksum = int(acts[(args.steps - 1) % n_rot].sum().item())

    # ---- end to end through the public API: host records -> actions ---------------------------------
    e2e = None
    if not args.no_e2e:
        agent = AgentVal(model, fe)
        host_recs = recs  # numpy records in pageable host memory, as a driver would hold them
        agent.act_batch(host_recs[0])
        for i in range(max(3, min(args.warmup, 10))):
            agent.act_batch(host_recs[i % n_rot])
        torch.cuda.synchronize()
        barrier()
        te0 = time.perf_counter()
        for i in range(args.steps):
            a = agent.act_batch(host_recs[i % n_rot])  # pack into pinned staging, H2D, kernel, D2H, sync
        torch.cuda.synchronize()
        te = max_over_ranks(time.perf_counter() - te0)
        barrier()
        assert a.shape == (B,)
        # the same work with two calls in flight (DecisionPipeline: what a lock-step driver over two groups of battles
        # does): every step still packs its records into pinned staging, moves them to the device, runs the fused
        # kernel with features materialised, and reads Q and actions back on the host
        from pokeai_b200 import DecisionPipeline, pinned_records
        depth = 3
        pipe = DecisionPipeline(agent, B, depth=depth)
        nw = max(3, min(args.warmup, 10))

        def timed(batches, pinned, n):
            torch.cuda.synchronize()
            barrier()
            t0 = time.perf_counter()
            for out in pipe.map((batches[i % n_rot] for i in range(n)), pinned=pinned):
                pass
            torch.cuda.synchronize()
            dt = max_over_ranks(time.perf_counter() - t0)
            barrier()
            return dt, out

        # (a) records in the caller's pageable memory: copied into pinned staging every step
        _, a2 = timed(host_recs, False, nw)
        assert np.array_equal(a2, agent.act_batch(host_recs[(nw - 1) % n_rot]))
        tpg, a2 = timed(host_recs, False, args.steps)
        # (b) records already in pinned host memory (a driver packs its requests straight into pinned_records()):
        # the kernel reads them over PCIe, no staging copy
        pin_recs = []
        for r in host_recs:
            p_ = pinned_records(B)
            p_.view(np.uint8)[...] = r.view(np.uint8)
            pin_recs.append(p_)
        _, a3 = timed(pin_recs, True, nw)
        assert np.array_equal(a3, agent.act_batch(host_recs[(nw - 1) % n_rot]))
        tp, a3 = timed(pin_recs, True, args.steps)
        assert a3.shape == (B,) and np.array_equal(a3, a2)
        pipe.close()
        e2e = {"value": B * world * args.steps / tp, "unit": UNIT, "h2d_bytes_per_step": B * 144,
               "d2h_bytes_per_step": B * A * 4 + B * 4, "ms_per_step": tp / args.steps * 1e3,
               "api": f"DecisionPipeline(agent, depth={depth}).map(batches, pinned=True) -> int32 actions per batch: every "
                      "step's packed records sit in pinned host memory and are read by the kernel over PCIe (the H2D "
                      "copy), features are materialised in HBM, Q and actions are written to pinned host memory (the "
                      "D2H) and the actions copied out; up to three steps in flight",
               "pageable_input_value": B * world * args.steps / tpg, "pageable_input_ms_per_step": tpg / args.steps * 1e3,
               "pageable_input_api": "the same with the records in ordinary host memory: one more 590 KB host copy into "
                                     "pinned staging per step",
               "sync_value": B * world * args.steps / te, "sync_ms_per_step": te / args.steps * 1e3,
               "sync_api": "AgentVal.act_batch(packed records): the same, one blocking call per step"}
    sampler.stop()

    cpu = None
    if rank == 0 and world == 1 and args.cpu_seconds > 0:
        cpu = run_cpu_baseline(recs[0], args, args.cpu_seconds)

    if rank == 0:
        peak, peak_src = peaks()
        ms_per_step = dev_ms / args.steps
        achieved = ALG_BYTES_PER_DECISION * B / (ms_per_step * 1e-3) / 1e9
        traffic = None
        tf = ROOT / "profiles" / "traffic.json"
        if tf.exists():
            try:
                traffic = json.loads(tf.read_text()).get(f"pk_decide_kernel@B={B}")
            except Exception:
                traffic = None
        line = {
            "metric": METRIC, "value": B * world * args.steps / (dev_ms * 1e-3), "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": dict(workload_config(args, world),
                           overlap="programmatic dependent launch: the record-only prologue of launch i+1 overlaps the "
                                   "tail of launch i (PK_PDL, on by default in the library)",
                           launch=(f"CUDA graph of {G} launches replayed {args.steps // G}x + {args.steps % G} plain launches"
                                   if graph is not None else "one Python-issued launch per step"),
                           l2=f"{n_rot} rotating input/output sets ({n_rot * feat_bytes / 1e6:.0f} MB of outputs) "
                              f"> 2x the 126 MB L2; no flush needed"),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "kernel": "pk_decide_kernel<2,9>",
                         "algorithmic_bytes_per_launch": ALG_BYTES_PER_DECISION * B, "peak_source": peak_src,
                         "launch_us": ms_per_step * 1e3},
            "e2e": e2e, "cpu_baseline": cpu, "clocks": clocks, "gpu_launches": args.steps,
            "checksum": checksum,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", 0))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    if args.impl == "reference":
        reference_arm(args, rank, world)
        return
    if world != args.gpus and rank == 0:
        print(f"bench.py: --gpus {args.gpus} but WORLD_SIZE={world}; launch with torch.distributed.run for N > 1",
              file=sys.stderr)
    b200_arm(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
