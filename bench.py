#!/usr/bin/env python
"""bench.py -- policy decisions/sec of the decision hot path (encode + MLP + masked select).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W

Workload = BASELINE.json configs[1]: batched inference over 4096 synthetic Gen-1 6v6 battle states per launch
(A = 9 action slots, C = 617 channels), default MLP 2 x 64, features and masks materialised in HBM, Q values and the
masked-argmax action written.  One "step" = one pass over one batch of 16 x 4096 = 65 536 states per GPU: 16
consecutive launches of the decision kernel (one CUDA-graph replay) rotating over four input/output sets, so that the
K timed steps last milliseconds, not microseconds.  `value` is decisions per second.  At N > 1 every rank runs the
same per-GPU batch on its own shard of states (battles are independent; no data-path collective) => weak scaling.

Prints ONE JSON line (rank 0).  Keys beyond the base contract:
  roofline      achieved HBM GB/s of the decision kernel = algorithmic bytes per launch (22 432 B per decision,
                SURVEY.md 8d, x 4096) / CUDA-event time per launch.  `roofline.secondary` carries the other BASELINE
                configurations, measured in the same process (bench_secondary.py): the 65 536-state shard, the DQN
                update with its gradient exchange (both routes + `parity_ok` at N > 1), the wide 4 x 2048 net on the
                tensor cores, 64 concurrent battles in lock step, and the reference's own eager-PyTorch GPU path.
  e2e           the same metric through the host-buffer public API, every step moving its records host -> device and
                its Q + actions device -> host: `value` = DecisionPipeline (three launches in flight, records in pinned
                host memory read by the kernel over PCIe); `pageable_input_value` = the same from ordinary host memory
                (one more host copy per launch); `sync_value` = one blocking AgentVal.act_batch call per launch.
  cpu_baseline  the reference's own CPU implementation of the path (oracle/_ref: the unmodified reference modules
                staged by build(); falls back to the port oracle/ref_port.py when absent) on this box's host cores.
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
LAUNCHES_PER_STEP = 16


def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=50)
    p.add_argument("--warmup", type=int, default=5)
    p.add_argument("--impl", default="b200", choices=["b200", "reference"])
    p.add_argument("--batch", type=int, default=4096, help="battle states per GPU per launch")
    p.add_argument("--layers", type=int, default=2)
    p.add_argument("--channels", type=int, default=64)
    p.add_argument("--cpu-seconds", type=float, default=12.0, help="CPU baseline budget (0 = skip)")
    p.add_argument("--no-e2e", action="store_true")
    p.add_argument("--no-secondary", action="store_true", help="skip the secondary legs (roofline.secondary)")
    p.add_argument("--skip-legs", default="", help="comma-separated secondary legs to skip")
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
    """identical for both arms (the driver compares the two `config` objects)"""
    G = LAUNCHES_PER_STEP
    return {
        "workload": f"Gen-1 6v6 batched inference, {args.batch} synthetic battle states per launch per GPU, "
                    f"default MLP {args.layers}x{args.channels}, features+mask+Q+action written (BASELINE configs[1]); "
                    f"one step = {G} consecutive launches = {G * args.batch} decisions per GPU",
        "states_per_launch": args.batch, "launches_per_step": G, "states_per_gpu_per_step": G * args.batch,
        "global_states_per_step": G * args.batch * world,
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


# ---- the reference's CPU implementation on the host cores (cpu_baseline and the --impl reference arm) ------------
def cpu_decide_factory(rec, n_layers, n_channels, threads):
    """Returns (step(n) -> decisions done, kind).  The ONLY place bench.py's CPU legs touch oracle/.
    kind "reference": the UNMODIFIED reference modules staged in oracle/_ref by build() (oracle/ref_stage.py) run
    exactly as the reference runs: a Python loop of FeatureExtractor.transform over observation objects, np.stack,
    Agent._calc_q_vector_batch (MLPModel on torch CPU fp32, all host threads), mask -> -inf, argmax
    (feature_extractor.py:40-52, agent.py:28-35).  kind "port": the restatement oracle/ref_port.py, pinned to the
    reference by tests/golden/, when oracle/_ref is absent."""
    import torch
    from oracle import ref_port, ref_stage
    from pokeai_b200.packed import unpack_observation
    from pokeai_b200.tables import load_tables
    T = load_tables()
    torch.set_num_threads(threads)
    obs = [unpack_observation(r, T) for r in rec]  # the reference receives objects; building them is not timed
    sd = ref_port.init_mlp_params(T.n_channels, n_layers, n_channels, seed=0)
    R = ref_stage.load()
    if R is not None:
        robs = [ref_stage.to_ref_obs(R, o) for o in obs]
        ref_step = ref_stage.decide_step(R, robs, PARTY_SIZE, sd, n_layers, n_channels)

        def step(n):
            return int(ref_step(n).shape[0])
        return step, "reference"

    def step(n):
        feats, masks = ref_port.transform_batch(obs[:n], PARTY_SIZE, T)
        with torch.no_grad():
            q = ref_port.mlp_forward(sd, torch.from_numpy(feats)).numpy()
        act = ref_port.masked_argmax(q, masks)
        return int(act.shape[0])

    return step, "port"


def host_threads():
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except Exception:
        return os.cpu_count() or 1


def run_cpu_baseline(rec, args, budget_s):
    threads = host_threads()
    step, kind = cpu_decide_factory(rec, args.layers, args.channels, threads)
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
    return {"value": n / med, "unit": UNIT, "cores": threads, "kind": kind,
            "sample": f"{n} of the {len(rec)} states of one launch, {len(times)} passes, median; encode loop is "
                      f"single-threaded Python as in the reference, torch forward uses {threads} threads"}


def reference_arm(args, rank, world):
    """--impl reference: the reference's own CPU implementation of the path on the host cores (oracle/_ref: the
    unmodified reference modules, staged by build(); the port oracle/ref_port.py only when that is absent)."""
    if rank != 0:
        return
    from pokeai_b200.packed import synth_packed
    rec = synth_packed(args.batch, seed=0, party_size=PARTY_SIZE, kind="full")
    threads = host_threads()
    step, kind = cpu_decide_factory(rec, args.layers, args.channels, threads)
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
    sample = (f"each step = the first {per_step} states of the workload's {LAUNCHES_PER_STEP * args.batch}; "
              f"Python encode loop (single-threaded, as the reference) + torch CPU fp32 forward on {threads} threads + "
              f"masked argmax")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, max(1, args.gpus)),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": kind, "sample": sample},
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

    # programmatic dependent launch is opt-in for the device-pointer API (include/pokeai_b200.h): the records of every
    # launch were uploaded above, none is produced by a preceding kernel, so the contract holds
    from pokeai_b200 import _native as N
    N.check(N.lib().pk_set_overlap_launches(1))

    def step(i):
        j = i % n_rot
        ops.decide(pk[j], A, prepared, nl, H, out_feat=feats[j], out_mask=masks[j], out_q=qs[j], out_action=acts[j])

    # One step = G consecutive launches (the rotation over the output sets included).  The G launches are captured in a
    # CUDA graph once and a step is one replay: same kernels, same buffers, same count, but no host work between
    # launches, so the number is the GPU's and not the Python launch loop's (which the NVML sampler thread slows
    # further through the GIL).  --no-graph issues the G launches of every step from Python.
    side = torch.cuda.Stream(device=dev)
    graph, G = None, LAUNCHES_PER_STEP
    sampler = ClockSampler(local_rank)
    sampler.start()
    warm = max(3, args.warmup)
    with torch.cuda.stream(side):
        for i in range(G):
            step(i)
        side.synchronize()
        if not args.no_graph:
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=side):
                for i in range(G):
                    step(i)

        def run_step():
            if graph is not None:
                graph.replay()
            else:
                for i in range(G):
                    step(i)

        for _ in range(warm):
            run_step()
        side.synchronize()
        barrier()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        side.synchronize()
        t0 = time.perf_counter()
        ev0.record(side)
        for _ in range(args.steps):
            run_step()
        ev1.record(side)
        side.synchronize()
        t1 = time.perf_counter()
    barrier()
    dev_ms = max_over_ranks(ev0.elapsed_time(ev1))
    clocks = sampler.summary(t0, t1)
    checksum = int(acts[(G - 1) % n_rot].sum().item())
    n_launch = args.steps * G  # launches inside the timed region

    # ---- end to end through the public API: host records -> actions ---------------------------------
    # The same steps (G launches of B states each) with every launch's records starting in HOST memory and its Q and
    # actions ending there.
    e2e = None
    if not args.no_e2e:
        agent = AgentVal(model, fe)
        host_recs = recs  # numpy records in pageable host memory, as a driver would hold them
        agent.act_batch(host_recs[0])
        n_e2e = args.steps * G               # launches of the pipelined legs: the same count as the device-timed region
        n_sync = min(n_e2e, 20 * G)          # the blocking leg is ~100 us per launch: bounded
        for i in range(max(3, min(warm, 10))):
            agent.act_batch(host_recs[i % n_rot])
        torch.cuda.synchronize()
        barrier()
        te0 = time.perf_counter()
        for i in range(n_sync):
            a = agent.act_batch(host_recs[i % n_rot])  # pageable records: H2D copy, kernel, results to pinned memory, sync
        torch.cuda.synchronize()
        te = max_over_ranks(time.perf_counter() - te0)
        barrier()
        assert a.shape == (B,)
        # the same work with several calls in flight (DecisionPipeline: what a lock-step driver over groups of battles
        # does): every launch still gets its records from host memory, runs the fused kernel with features
        # materialised, and hands Q and actions back on the host
        from pokeai_b200 import DecisionPipeline, pinned_records
        depth = 3
        pipe = DecisionPipeline(agent, B, depth=depth)
        nw = max(3, min(warm, 10))

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

        # (a) records in the caller's pageable memory: copied into pinned staging every launch
        _, a2 = timed(host_recs, False, nw)
        assert np.array_equal(a2, agent.act_batch(host_recs[(nw - 1) % n_rot]))
        tpg, a2 = timed(host_recs, False, n_e2e)
        # (b) records already in pinned host memory (a driver packs its requests straight into pinned_records()):
        # the kernel reads them over PCIe, no staging copy
        pin_recs = []
        for r in host_recs:
            p_ = pinned_records(B)
            p_.view(np.uint8)[...] = r.view(np.uint8)
            pin_recs.append(p_)
        _, a3 = timed(pin_recs, True, nw)
        assert np.array_equal(a3, agent.act_batch(host_recs[(nw - 1) % n_rot]))
        tp, a3 = timed(pin_recs, True, n_e2e)
        assert a3.shape == (B,) and np.array_equal(a3, a2)
        pipe.close()
        e2e = {"value": B * world * n_e2e / tp, "unit": UNIT, "h2d_bytes_per_step": G * B * 144,
               "d2h_bytes_per_step": G * (B * A * 4 + B * 4), "ms_per_step": tp / args.steps * 1e3,
               "api": f"DecisionPipeline(agent, depth={depth}).map(batches, pinned=True) -> int32 actions per batch of "
                      f"{B}: every launch's packed records sit in pinned host memory and are read by the kernel over PCIe "
                      "(the H2D copy), features are materialised in HBM, Q and actions are written to pinned host memory "
                      f"(the D2H) and the actions copied out; up to {depth} launches in flight; {G} launches per step",
               "pageable_input_value": B * world * n_e2e / tpg, "pageable_input_ms_per_step": tpg / args.steps * 1e3,
               "pageable_input_api": "the same with the records in ordinary host memory: one more 590 KB host copy into "
                                     "pinned staging per launch",
               "sync_value": B * world * n_sync / te, "sync_ms_per_launch": te / n_sync * 1e3,
               "sync_api": f"AgentVal.act_batch(packed records in pageable host memory): one blocking call per launch "
                           f"({n_sync} calls timed; 61 us per call with the records in pinned_records(), tools/prof_act_batch.py)"}
    sampler.stop()

    # ---- the other BASELINE configurations, same process (bench_secondary.py) -> roofline.secondary ------------------
    secondary = None
    if not args.no_secondary:
        from types import SimpleNamespace
        import bench_secondary
        ctx = SimpleNamespace(dev=dev, rank=rank, world=world, dist=dist if world > 1 else None, barrier=barrier,
                              max_over_ranks=max_over_ranks, prepared=prepared, model=model, fe=fe)
        secondary = bench_secondary.run_all(ctx, skip=tuple(x for x in args.skip_legs.split(",") if x))

    cpu = None
    if rank == 0 and world == 1 and args.cpu_seconds > 0:
        cpu = run_cpu_baseline(recs[0], args, args.cpu_seconds)
        if secondary:
            gref = [r for r in secondary if r.get("leg") == "gpu_reference" and "device_resident_per_s" in r]
            if gref:
                cpu["gpu_reference"] = {"what": "the reference's eager-PyTorch GPU path on pre-encoded features "
                                                "(details in roofline.secondary)",
                                        "default_mlp_decisions_per_s": gref[0]["device_resident_per_s"],
                                        "wide_mlp_decisions_per_s": gref[-1]["device_resident_per_s"], "unit": UNIT}

    if rank == 0:
        peak, peak_src = peaks()
        ms_per_step = dev_ms / args.steps
        launch_ms = ms_per_step / G
        achieved = ALG_BYTES_PER_DECISION * B / (launch_ms * 1e-3) / 1e9
        traffic = None
        tf = ROOT / "profiles" / "traffic.json"
        if tf.exists():
            try:
                tj = json.loads(tf.read_text())
                traffic = tj.get(f"pk_decide_kernel@B={B}")
            except Exception:
                traffic = None
        line = {
            "metric": METRIC, "value": B * world * n_launch / (dev_ms * 1e-3), "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": warm, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args, world),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic,
                         "kernel": ("pk_decide_tc_kernel<9>" if (os.environ.get("PK_DECIDE_TC", "0") not in ("", "0")
                                                                  and nl >= 2 and 32 < H <= 64)
                                    else "pk_decide_kernel<2,9,512,0>"),
                         "algorithmic_bytes_per_launch": ALG_BYTES_PER_DECISION * B, "peak_source": peak_src,
                         "launch_us": launch_ms * 1e3, "launches_timed": n_launch,
                         "overlap": "programmatic dependent launch, opted into with pk_set_overlap_launches(1): the "
                                    "record-only prologue of launch i+1 overlaps the tail of launch i",
                         "launch": (f"one CUDA-graph replay of {G} launches per step" if graph is not None
                                    else f"{G} Python-issued launches per step"),
                         "l2": f"{n_rot} rotating input/output sets ({n_rot * feat_bytes / 1e6:.0f} MB of outputs) "
                               f"> 2x the 126 MB L2; no flush needed",
                         "secondary": secondary},
            "e2e": e2e, "cpu_baseline": cpu, "clocks": clocks, "gpu_launches": n_launch,
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
