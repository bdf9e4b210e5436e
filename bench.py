#!/usr/bin/env python
"""bench.py -- throughput of the AIrcraft hot path on B200 (driver contract: one JSON line).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl native|reference] [--workload c4|c2]

Metric (BASELINE.json): surrogate-RK4 state-steps/s AND Jacobian evals/s at 1/2/4/8 B200 next to the CPU path on the
box's host cores.  Everything the metric names is a compact TOP-LEVEL block of the one JSON line:

  value / e2e / roofline   main line: the per-GPU shard of BASELINE config 4 (weak scaling) -- the sampling-MPC sweep
        "1M control-sequence rollouts x horizon 200 sharded over 8 B200" = 125 000 rollouts x 200 RK4 steps per GPU
        from one shared initial state; controls SAMPLED IN THE KERNEL around one nominal sequence, fused loss,
        warp-shuffle arg-min, winner reproduced from checkpoints; N>1: all-gather of the 16 kB winner records +
        device-side selection.  At N=8 the job is exactly config 4.  (c4 is the largest single-GPU state-step config
        and the one the 1e10 target is quoted on; config 2 fits one GPU too and is reported under "c2".)
  sustained                the same step back to back for >= 2 s with the clock sampler running (>= 10 samples)
  config4_strong           config 4 as BASELINE words it: 1 000 000 rollouts TOTAL over N GPUs (strong scaling)
  c2                       BASELINE config 2: 4096 rollouts x 200 from trim states on ONE GPU (rank 0), latency case
  jacobian                 Jacobian evals/s aggregated over the N ranks (max time over ranks): BASELINE config 3
        (65 536 x 100 nodes PER GPU, weak) with roofline in GB/s and executed FLOP/s, BASELINE config 5 exactly
        (8192 windows x 50 nodes TOTAL over N GPUs), e2e through airc_linearize_host (pinned host buffers), CPU baseline
  extra                    config 1, explicit-control shard, Hessian blocks, the reference's 298-node MPC blocks

`--impl reference` times the CPU restatement of the reference path (oracle/cpu_ref.cpp, float64, OpenMP over all host
cores; the reference's own CasADi evaluation cannot run on this box) on the same workload, each step a bounded sample.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

F_STEP = 3953          # algorithmic FLOPs per RK4 state-step (minimal form)
B_JAC = 1172           # algorithmic HBM bytes per Jacobian eval: x,u in; x+, A, B out (FP32)
F_JAC = 28000          # estimated algorithmic FLOPs per Jacobian eval
DT = 0.01
H_ROLL = 200
C4_TOTAL, C4_GPUS = 1_000_000, 8
FP32_NOMINAL_TFLOPS = 148 * 128 * 2 * 1.965e9 / 1e12     # 74.4


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--workload", default="c4", choices=["c4", "c2"])
    ap.add_argument("--no-extras", action="store_true", help="skip the c2/c3/c5 side measurements")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="CPU-baseline sample budget")
    ap.add_argument("--per-gpu", type=int, default=0, help="override rollouts per GPU (debug)")
    return ap.parse_args()


# ----------------------------------------------------------------------------- clocks
class ClockSampler:
    """nvidia-smi sampled every 200 ms DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                 "--format=csv,noheader,nounits", "-lms", "200"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def mark(self):
        return time.time()

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(2)
            except Exception:
                self.proc.kill()

    def summary(self, windows):
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for t, line in self.rows:
            if not any(a - 0.05 <= t <= b + 0.25 for a, b in windows):
                continue
            f = [c.strip() for c in line.split(",")]
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); pw.append(float(f[2]))
            except Exception:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)),
                "power_w_max": float(max(pw)), "reasons": sorted(reasons), "samples": len(sm)}


# ----------------------------------------------------------------------------- CPU legs
def _cpu_ref(threads: int = 0):
    """The float64 C++ restatement (oracle/cpu_ref.cpp).  bench.py may execute oracle/ only here:
    the cpu_baseline leg and the --impl reference arm."""
    from aircraft_b200.config import AircraftConfiguration, TRIM_COM
    from aircraft_b200.params import load_nn_blob, make_params
    from oracle.cpu_ref import CpuRef
    p = make_params(AircraftConfiguration.from_json(), com=TRIM_COM)
    return CpuRef(p, load_nn_blob(), threads=threads)


def _cpu_sample_inputs(n: int, horizon: int):
    from aircraft_b200.workloads import make_control_sequences, make_states
    x0 = np.repeat(make_states(1, seed=4, level=True), n, axis=0)
    U = make_control_sequences(n, horizon, seed=4)
    return x0, U


def sweep_inputs():
    """The shared initial state and nominal control sequence of the sweep (host, float32/64)."""
    from aircraft_b200.workloads import make_control_sequences, make_states
    x0 = make_states(1, seed=4, level=True)[0]
    Un = make_control_sequences(1, H_ROLL, seed=4, sigma=0.05)[0].astype(np.float32)
    return x0, Un


def sweep_sampler(seed: int = 2025, offset: int = 0):
    from aircraft_b200.params import make_sweep_params
    return make_sweep_params(seed=seed, index_offset=offset)


def cpu_sweep_once(ref, n: int):
    """One CPU pass of the same job: sample n control sequences with the same generator, roll them
    out in float64, loss, (min, arg-min)."""
    from aircraft_b200.params import make_cost_params
    x0, Un = sweep_inputs()
    t = time.perf_counter()
    ref.sweep(x0, Un, DT, n, sweep_sampler(), make_cost_params(), normalise=True, return_costs=False)
    return time.perf_counter() - t


def cpu_rollout_rate(ref, seconds: float, horizon: int = H_ROLL):
    """state-steps/s of the CPU restatement on a bounded sample of the c4 workload: calibrate on
    one trajectory per thread, then size the sample to ~`seconds` of work."""
    n0 = max(ref.threads, 1)
    per_traj = max(cpu_sweep_once(ref, n0), 1e-4)      # one trajectory per thread in parallel
    n = n0 * int(max(1, min(4096, seconds / per_traj)))
    el = cpu_sweep_once(ref, n)
    return n * horizon / el, n, el


def cpu_jacobian_rate(ref, seconds: float):
    from aircraft_b200.workloads import make_controls, make_states
    n0 = max(ref.threads, 1) * 16
    x, u = make_states(n0, seed=5, perturb_norm=1e-3), make_controls(n0, seed=5)
    t = time.perf_counter(); ref.linearize(x, u, DT, normalise=True); t0 = max(time.perf_counter() - t, 1e-4)
    n = int(max(n0, min(2_000_000, n0 * seconds / t0)))
    x, u = make_states(n, seed=5, perturb_norm=1e-3), make_controls(n, seed=5)
    t = time.perf_counter(); ref.linearize(x, u, DT, normalise=True); el = time.perf_counter() - t
    return n / el, n, el


def run_reference(args):
    """--impl reference: the CPU restatement of the reference path, all host threads, same metric
    and config; each step is a bounded sample of the workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    ref = _cpu_ref()
    per_gpu = args.per_gpu or (C4_TOTAL // C4_GPUS if args.workload == "c4" else 4096)
    # size one step to ~2 s so that K+W steps end within a few minutes
    rate, _, _ = cpu_rollout_rate(ref, 1.0)
    n = int(max(ref.threads, min(per_gpu, rate * 2.0 / H_ROLL)))
    for _ in range(args.warmup):
        cpu_sweep_once(ref, n)
    el = 0.0
    for _ in range(args.steps):
        el += cpu_sweep_once(ref, n)
    value = n * H_ROLL * args.steps / el
    jrate, jn, _ = cpu_jacobian_rate(ref, 3.0)
    line = {
        "impl": "reference", "metric": "surrogate RK4 state-steps/s", "value": value,
        "unit": "state-steps/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": el / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args.workload, per_gpu, args.gpus),
        "cpu_baseline": {"value": value, "unit": "state-steps/s", "cores": ref.threads, "kind": "port",
                         "sample": f"{n} sampled rollouts x {H_ROLL} steps per step (of {per_gpu} per GPU), same "
                                   "sampler + loss + arg-min, oracle/cpu_ref.cpp float64 OpenMP; the "
                                   "reference's CasADi VM cannot run here"},
        "jacobian": {"value": jrate, "unit": "jacobian-evals/s", "sample": f"{jn} nodes, 20-direction forward mode"},
        "e2e": {"value": value, "unit": "state-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def workload_config(workload: str, per_gpu: int, gpus: int):
    if workload == "c4":
        name = (f"BASELINE config 4 shard: sampling-MPC sweep, {per_gpu} control-sequence rollouts x "
                f"horizon {H_ROLL} per GPU from one shared state ({per_gpu * gpus} total; "
                f"1M total at 8 GPUs), controls sampled in the kernel, fused cost + arg-min + "
                f"best-trajectory record exchange")
    else:
        name = f"BASELINE config 2: {per_gpu} rollouts x horizon {H_ROLL} from trim states per GPU"
    return {"workload": name, "rollouts_per_gpu": per_gpu, "horizon": H_ROLL, "dt": DT,
            "substeps": 1, "normalise": True, "layout": "soa", "parallelism": f"dp{gpus}",
            "l2_policy": "no reusable inputs: each launch reads 5.7 kB and streams > 126 MB of checkpoint/"
                         "slice records through L2, which evicts everything between timed iterations"
            if workload == "c4" else "L2 flushed between timed iterations (256 MB write)"}


# ----------------------------------------------------------------------------- GPU legs
def device_control_sequences(torch, n: int, horizon: int, seed: int, device):
    """[H,7,n] FP32 random-walk control sequences generated on the device (synthetic data; same
    distribution family as workloads.make_control_sequences, clipped to the actuator limits)."""
    g = torch.Generator(device=device).manual_seed(20250101 + seed)
    lo = torch.tensor([-5.0, -5.0, -5.0, 0.0, 0.0, 0.0, 0.0], device=device).view(1, 7, 1)
    hi = torch.tensor([5.0, 5.0, 5.0, 0.0, 0.0, 0.0, 1.0], device=device).view(1, 7, 1)
    sg = torch.tensor([0.25, 0.25, 0.25, 0.0, 0.0, 0.0, 0.025], device=device).view(1, 7, 1)
    # a handful of launches (not 4 per step of the horizon), so that an ncu launch list of the
    # bench command is dominated by the product's kernels
    U = torch.randn((horizon, 7, n), generator=g, device=device, dtype=torch.float32)
    U.mul_(sg)
    U[0] = (lo + (hi - lo) * torch.rand((1, 7, n), generator=g, device=device))[0]
    torch.cumsum(U, dim=0, out=U)
    return torch.minimum(torch.maximum(U, lo), hi, out=U)


def time_launches(torch, fn, steps: int, warmup: int, flush=None):
    """Average device time of fn() in ms: CUDA events on the current stream around every call,
    optional L2 flush before each."""
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    evs = []
    for _ in range(steps):
        if flush is not None:
            flush.add_(1.0)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        evs.append((e0, e1))
    torch.cuda.synchronize()
    return sum(a.elapsed_time(b) for a, b in evs) / steps


def run_native(args):
    import torch
    import torch.distributed as dist
    from aircraft_b200 import Aircraft
    from aircraft_b200.config import TRIM_COM
    from aircraft_b200.workloads import make_states, trim_states

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; aircraft_b200 has no CPU path (use --impl reference)")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        # NCCL writes its "NCCL version ..." banner to STDOUT at NCCL_DEBUG=VERSION (this image's setting);
        # stdout must carry exactly one JSON line, so the banner level is lowered unless the caller asked for more
        if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
            os.environ["NCCL_DEBUG"] = "WARN"
        dist.init_process_group("nccl", device_id=dev)
    if world != args.gpus and rank == 0:
        print(f"bench.py: --gpus {args.gpus} but WORLD_SIZE={world}; using WORLD_SIZE", file=sys.stderr)

    ac = Aircraft.from_problem(com=TRIM_COM, device=local, physical_integration_substeps=1, normalise=True)
    per_gpu = args.per_gpu or (C4_TOTAL // C4_GPUS if args.workload == "c4" else 4096)
    total = per_gpu * world
    H, K, W = H_ROLL, args.steps, max(args.warmup, 3)

    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    windows = []

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    if args.workload == "c2":
        line = run_c2(torch, dist, ac, dev, rank, world, per_gpu, K, W, clocks, windows, barrier)
    else:
        line = run_c4(torch, dist, ac, dev, rank, world, per_gpu, total, K, W, clocks, windows, barrier)

    peak = line["roofline"]["peak"]

    def rank_max(v: float) -> float:
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    if not args.no_extras:
        # ---- BASELINE config 2 (one GPU): rank 0 measures, the others wait
        if rank == 0 and args.workload != "c2":
            line["c2"] = measure_c2(torch, ac, dev, peak, max(5, min(K, 20)))
        barrier()
        # ---- Jacobian evals/s over all ranks: config 3 per GPU (weak), config 5 exactly (strong), e2e
        line["jacobian"] = measure_jacobian(torch, ac, dev, rank, world, max(5, min(K, 20)), peak, rank_max, barrier,
                                            clocks, windows)
        barrier()
        if rank == 0:
            line["extra"] = side_extras(torch, ac, dev, args)
    if rank == 0 and world == 1:
        ref = _cpu_ref()
        rate, n, sec = cpu_rollout_rate(ref, args.cpu_seconds)
        line["cpu_baseline"] = {"value": rate, "unit": "state-steps/s", "cores": ref.threads, "kind": "port",
                                "sample": f"{n} sampled rollouts x {H} steps ({sec:.1f} s) of the same workload "
                                          "(same sampler, loss and arg-min), oracle/cpu_ref.cpp float64 OpenMP "
                                          "(CasADi cannot run on this box)"}
        if "jacobian" in line:
            jr, jn, js = cpu_jacobian_rate(ref, 4.0)
            line["jacobian"]["cpu_baseline"] = {"value": jr, "unit": "jacobian-evals/s", "cores": ref.threads,
                                                "kind": "port", "sample": f"{jn} nodes ({js:.1f} s), 20-direction forward mode"}
    if rank == 0:
        clocks.stop()
        line["clocks"] = clocks.summary(windows)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def fp32_peak(ac):
    """FP32 FFMA peak of this device, measured now (MEASURED_PEAKS.json has no CUDA-core entry)."""
    import ctypes as C
    tf, ms = C.c_double(), C.c_double()
    ac._lib.airc_measure_fp32_peak(ac.handle, 5, C.byref(tf), C.byref(ms))
    return float(tf.value)


def csrc_sha():
    import hashlib
    h = hashlib.sha256()
    for p in sorted((ROOT / "aircraft_b200" / "csrc").glob("*")):
        h.update(p.name.encode())
        h.update(p.read_bytes())
    return h.hexdigest()[:16]


def kernel_audit(name: str):
    """One entry of profiles/r2_kernel_audit.json (written by profiles/ncu_audit.py from an `ncu --set full` capture
    of this workload): DRAM bytes and EXECUTED FP32 flops per launch of the kernel.  ncu cannot run inside the bench,
    so the entry carries the hash of the CUDA sources it was captured on and is reported as stale when they changed."""
    try:
        e = json.loads((ROOT / "profiles" / "r2_kernel_audit.json").read_text())[name]
        e = dict(e)
        e["stale"] = e.get("csrc_sha") != csrc_sha()
        return e
    except Exception:
        return None


def rollout_roofline(kernel, tflops, peak, steps_per_launch, ms_kernel, ms_step, note="", audit=None):
    r = {"kernel": kernel, "bound": "fp32", "achieved": tflops, "peak": peak, "unit": "TFLOP/s",
         "frac": tflops / peak,
         "peak_source": "FFMA probe measured in this run (airc_measure_fp32_peak); MEASURED_PEAKS.json has no FP32 "
                        "CUDA-core entry",
         "peak_nominal": FP32_NOMINAL_TFLOPS, "frac_of_nominal": tflops / FP32_NOMINAL_TFLOPS,
         "flops_per_state_step": F_STEP, "state_steps_per_launch": steps_per_launch,
         "kernel_ms": ms_kernel, "kernel_share_of_step": ms_kernel / ms_step, "traffic": None, "note": note}
    if audit and abs(audit["units_per_launch"] - steps_per_launch) < 0.5:
        ex = audit["executed_fp32_flops_per_unit"]
        r.update({"traffic": audit["dram_bytes_per_launch"],
                  "traffic_source": f"profiles/{audit['capture']} via profiles/r2_kernel_audit.json (ncu --set full of this "
                                    f"workload; CUDA sources {'CHANGED since' if audit['stale'] else 'unchanged since'} "
                                    f"the capture); the kernel is FP32-bound, the traffic is checkpoint / slice records",
                  "executed_flops_per_state_step": ex,
                  "achieved_executed": tflops * ex / F_STEP, "frac_executed": tflops * ex / F_STEP / peak})
    return r


def run_c4(torch, dist, ac, dev, rank, world, per_gpu, total, K, W, clocks, windows, barrier):
    """BASELINE config 4 (shard): sampling sweep with in-kernel control sampling."""
    from aircraft_b200.params import make_cost_params
    from aircraft_b200.sharding import sweep_sampled
    H = H_ROLL
    x0_np, Un_np = sweep_inputs()
    x0_h = torch.tensor(x0_np, dtype=torch.float32).pin_memory()
    Un_h = torch.tensor(Un_np, dtype=torch.float32).pin_memory()
    x0_d, Un_d = x0_h.to(dev), Un_h.to(dev)
    sp, cp = sweep_sampler(), make_cost_params()
    ws = None
    torch.cuda.synchronize()

    # ---- `value`: inputs resident in HBM; whole step = sweep (prep + time-sliced rollout/cost/arg-min
    #      kernel + winner re-roll) and, for N>1, the all-gather of records + device-side selection
    roll_events = []

    def step(record: bool):
        nonlocal ws
        if record:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
        if world > 1:
            out = sweep_sampled(ac, x0_d, Un_d, DT, total, sampler=sp, cost=cp, workspace=ws, on_local=(
                (lambda: (e1.record(), roll_events.append((e0, e1)))) if record else None))
        else:
            out = ac.sweep(x0_d, Un_d, DT, per_gpu, sampler=sp, cost=cp, workspace=ws)
            if record:
                e1.record(); roll_events.append((e0, e1))
        ws = out.workspace
        return out

    # The timed step is ONE CUDA-graph launch per step (sweep kernels + all-gather + selection captured once,
    # aircraft_b200.sharding.SweepPlan); the kernels' own duration is measured afterwards, eagerly, with events
    # around the airc_sweep call.  If capture is not possible the eager step is timed instead (and said so).
    plan, graph_note = None, "one CUDA graph launch per step (SweepPlan: airc_sweep + all-gather + airc_select_record)"
    try:
        from aircraft_b200.sharding import SweepPlan
        plan = SweepPlan(ac, H, DT, total, sampler=sp, cost=cp)
        plan.x0.copy_(x0_d)
        plan.U_nom.copy_(Un_d)
    except Exception as e:                                   # noqa: BLE001 -- report, do not hide
        plan, graph_note = None, f"eager launches (graph capture failed: {type(e).__name__}: {e})"

    def timed_step():
        return plan.replay() if plan is not None else step(False)

    for _ in range(W):
        out = timed_step()
    barrier()
    t_mark = clocks.mark()
    s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s0.record()
    for _ in range(K):
        out = timed_step()
    s1.record()
    barrier()
    windows.append((t_mark, clocks.mark()))
    ms_total = s0.elapsed_time(s1)
    ms_by_rank = [ms_total / K]
    if world > 1:
        t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
        every = torch.empty(world, dtype=torch.float64, device=dev)
        dist.all_gather_into_tensor(every, t)
        ms_by_rank = [float(v) / K for v in every.tolist()]     # every rank's own device time: the spread between GPUs
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
    # kernel duration: the same sweep launched eagerly, CUDA events around airc_sweep (stream = torch's current)
    for _ in range(2):
        step(False)
    barrier()
    for _ in range(K):
        step(True)
    barrier()
    torch.cuda.synchronize()
    ms_step = ms_total / K
    value = total * H / (ms_step * 1e-3)
    launches = K * (3 + (1 if world > 1 else 0))      # prep, sweep2, winner2 (+ select_record)
    ms_roll = sum(a.elapsed_time(b) for a, b in roll_events) / K
    kernel_ms_by_rank = [ms_roll]
    if world > 1:   # every GPU's own kernel time (no exchange inside): the step waits for the slowest one
        every = torch.empty(world, dtype=torch.float64, device=dev)
        dist.all_gather_into_tensor(every, torch.tensor([ms_roll], dtype=torch.float64, device=dev))
        kernel_ms_by_rank = [float(v) for v in every.tolist()]
    roll_tflops = per_gpu * H * F_STEP / (ms_roll * 1e-3) / 1e12
    peak = fp32_peak(ac)

    # ---- `e2e`: host buffers in, host record out, every step.  N=1: the C-ABI host entry point
    #      (airc_sweep_host); N>1: pinned host -> device copies, sweep, exchange, record -> host.
    h2d = x0_h.numel() * 4 + Un_h.numel() * 4
    d2h = int(out.raw.numel()) * 4
    Ke = max(3, min(K, 20))
    x0_pin, Un_pin = x0_h.numpy(), Un_h.numpy()

    def e2e_step():
        if world == 1:
            r = ac.sweep(x0_pin, Un_pin, DT, per_gpu, sampler=sp, cost=cp)
            return r.min_cost
        xd, ud = x0_h.to(dev, non_blocking=True), Un_h.to(dev, non_blocking=True)
        r = sweep_sampled(ac, xd, ud, DT, total, sampler=sp, cost=cp, workspace=ws)
        return float(r.raw.cpu()[0])

    for _ in range(2):
        e2e_step()
    barrier()
    t_mark = clocks.mark()
    t0 = time.perf_counter()
    for _ in range(Ke):
        best = e2e_step()
    torch.cuda.synchronize()
    el = time.perf_counter() - t0
    windows.append((t_mark, clocks.mark()))
    if world > 1:
        t = torch.tensor([el], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        el = float(t.item())
    e2e_value = total * H * Ke / el

    # ---- sustained: the same step back to back for >= 2 s of device time, clock sampler running
    n_sus = int(max(50, min(4000, 2200.0 / ms_step)))
    barrier()
    t_mark = clocks.mark()
    s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s0.record()
    for _ in range(n_sus):
        timed_step()
    s1.record()
    barrier()
    sus_window = (t_mark, clocks.mark())
    windows.append(sus_window)
    ms_sus = s0.elapsed_time(s1)
    if world > 1:
        t = torch.tensor([ms_sus], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_sus = float(t.item())
    sustained = {"launches": n_sus, "seconds": ms_sus * 1e-3, "ms_per_step": ms_sus / n_sus,
                 "value": total * H / (ms_sus / n_sus * 1e-3), "unit": "state-steps/s",
                 "tflops_per_gpu": per_gpu * H * F_STEP / (ms_sus / n_sus * 1e-3) / 1e12,
                 "frac_of_fp32_peak": per_gpu * H * F_STEP / (ms_sus / n_sus * 1e-3) / 1e12 / peak,
                 "clocks": clocks.summary([sus_window]) if rank == 0 else None}

    # ---- config 4 as BASELINE words it: 1 000 000 rollouts in total over the N GPUs (strong scaling)
    strong = None
    if (per_gpu, H) == (C4_TOTAL // C4_GPUS, H_ROLL):
        ws_s = None

        def sstep():
            nonlocal ws_s
            if world > 1:
                o = sweep_sampled(ac, x0_d, Un_d, DT, C4_TOTAL, sampler=sp, cost=cp, workspace=ws_s)
            else:
                o = ac.sweep(x0_d, Un_d, DT, C4_TOTAL, sampler=sp, cost=cp, workspace=ws_s)
            ws_s = o.workspace
            return o

        for _ in range(2):
            so = sstep()
        barrier()
        Ks = max(5, min(K, 10))
        q0, q1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        q0.record()
        for _ in range(Ks):
            so = sstep()
        q1.record()
        barrier()
        ms_s = q0.elapsed_time(q1)
        if world > 1:
            t = torch.tensor([ms_s], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms_s = float(t.item())
        ms_s /= Ks
        strong = {"workload": f"BASELINE config 4 exactly: {C4_TOTAL} rollouts x {H} in total, {C4_TOTAL // world} per GPU",
                  "scaling": "strong", "n_gpus": world, "steps": Ks, "ms_per_step": ms_s,
                  "value": C4_TOTAL * H / (ms_s * 1e-3), "unit": "state-steps/s",
                  "tflops_per_gpu": C4_TOTAL / world * H * F_STEP / (ms_s * 1e-3) / 1e12,
                  "frac_of_fp32_peak": C4_TOTAL / world * H * F_STEP / (ms_s * 1e-3) / 1e12 / peak,
                  "best": {"min_cost": float(so.min_cost), "argmin": int(so.argmin)}}
        del ws_s

    return {
        "metric": "surrogate RK4 state-steps/s", "value": value, "unit": "state-steps/s",
        "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms_step, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "ms_per_step_by_rank": ms_by_rank, "kernel_ms_by_rank": kernel_ms_by_rank,
        "config": workload_config("c4", per_gpu, world),
        "sustained": sustained, "config4_strong": strong,
        "e2e": {"value": e2e_value, "unit": "state-steps/s", "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": d2h, "steps": Ke,
                "path": "airc_sweep_host (C ABI, host buffers in, winner record out)" if world == 1 else
                        "pinned host -> device, airc_sweep, NCCL all-gather of records, airc_select_record, record -> host"},
        "gpu_launches": launches, "launch_mode": graph_note,
        "best": {"min_cost": float(out.min_cost), "argmin": int(out.argmin)},
        "roofline": rollout_roofline("sweep2tc_kernel<ckpt>: network layers on tcgen05 (TF32 split in three, 18 432 tensor flop per "
                                     "state-step = 0.18 of the TF32 peak), the rest on FFMA2 (+ sweep2_prep, sweep_winner2: "
                                     "< 2 % of the call)",
                                     roll_tflops, peak, per_gpu * H, ms_roll, ms_step,
                                     "kernel_ms = CUDA events around the airc_sweep call (three launches)",
                                     audit=kernel_audit("sweep2tc")),
    }


def run_c2(torch, dist, ac, dev, rank, world, per_gpu, K, W, clocks, windows, barrier):
    """BASELINE config 2 as the main line (--workload c2): explicit controls from trim states."""
    from aircraft_b200.workloads import trim_states
    H = H_ROLL
    xs, _ = trim_states(ac, per_gpu, seed=rank)
    x0 = torch.tensor(xs.T, dtype=torch.float32, device=dev).contiguous()
    U = device_control_sequences(torch, per_gpu, H, seed=100 + rank, device=dev)
    flush = torch.zeros(64 * 1024 * 1024, dtype=torch.float32, device=dev)
    ms_roll = time_launches(torch, lambda: ac.rollout(x0, U, DT, layout="soa", return_final=False, cost=True), K, W, flush)
    barrier()
    if world > 1:
        t = torch.tensor([ms_roll], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_roll = float(t.item())
    x0_h = torch.empty((13, per_gpu), dtype=torch.float32).pin_memory(); x0_h.copy_(x0)
    U_h = torch.empty((H, 7, per_gpu), dtype=torch.float32).pin_memory(); U_h.copy_(U)
    x0_np, U_np = x0_h.numpy(), U_h.numpy()
    for _ in range(2):
        ac.rollout(x0_np, U_np, DT, layout="soa", return_final=False, cost=True)
    t0 = time.perf_counter()
    Ke = max(3, min(K, 10))
    for _ in range(Ke):
        ac.rollout(x0_np, U_np, DT, layout="soa", return_final=False, cost=True)
    el = time.perf_counter() - t0
    tfl = per_gpu * H * F_STEP / (ms_roll * 1e-3) / 1e12
    return {
        "metric": "surrogate RK4 state-steps/s", "value": per_gpu * world * H / (ms_roll * 1e-3),
        "unit": "state-steps/s", "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms_roll,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config("c2", per_gpu, world),
        "e2e": {"value": per_gpu * world * H * Ke / el, "unit": "state-steps/s",
                "h2d_bytes_per_step": x0_np.nbytes + U_np.nbytes, "d2h_bytes_per_step": 4 * per_gpu, "steps": Ke,
                "path": "airc_rollout_host (C ABI, pinned host buffers)"},
        "gpu_launches": K,
        "roofline": rollout_roofline("rollout_kernel<cost>", tfl, fp32_peak(ac), per_gpu * H, ms_roll, ms_roll),
    }


def measured_peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            return json.load(open(p)), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}, "fallback (B200_PROFILING.md)"


def measure_c2(torch, ac, dev, peak, K):
    """BASELINE config 2 -- 4096 rollouts x horizon 200 from random trim states on ONE GPU, explicit controls, fused
    cost: the latency case (4096 trajectories cannot fill 148 SMs with one trajectory per thread)."""
    from aircraft_b200.workloads import trim_states
    xs, _ = trim_states(ac, 4096, seed=0)
    x0 = torch.tensor(xs.T, dtype=torch.float32, device=dev).contiguous()
    U = device_control_sequences(torch, 4096, H_ROLL, seed=200, device=dev)
    flush = torch.zeros(64 * 1024 * 1024, dtype=torch.float32, device=dev)   # 256 MB > 126 MB L2
    ms_default = time_launches(torch, lambda: ac.rollout(x0, U, DT, layout="soa", cost=True), K, 3, flush)
    ms_lat = time_launches(torch, lambda: ac.rollout(x0, U, DT, layout="soa", cost=True, cooperative=True), K, 3, flush)
    ms = min(ms_default, ms_lat)
    tfl = 4096 * H_ROLL * F_STEP / (ms * 1e-3) / 1e12
    return {"workload": "BASELINE config 2: 4096 rollouts x horizon 200 from trim states, 1 GPU, U[H,7,B] explicit, fused cost",
            "ms_per_launch": ms, "value": 4096 * H_ROLL / (ms * 1e-3), "unit": "state-steps/s",
            "kernel": "latency mode: 8 lanes per trajectory (the default dispatch at this batch size)",
            "ms_default_dispatch": ms_default, "ms_latency_flag": ms_lat,
            "roofline": {"bound": "fp32 (latency-bound: 4096 trajectories on 148 SMs)", "achieved": tfl, "peak": peak,
                         "unit": "TFLOP/s", "frac": tfl / peak},
            "l2": "flushed between launches"}


def measure_jacobian(torch, ac, dev, rank, world, K, fp32_peak_tflops, rank_max, barrier, clocks, windows):
    """Jacobian evals/s over all N ranks (no collective on the data path; time = max over ranks)."""
    import ctypes as C
    from aircraft_b200 import _cabi
    from aircraft_b200.workloads import make_controls, make_states
    peaks, peak_src = measured_peaks()
    out = {"unit": "jacobian-evals/s", "n_gpus": world, "bytes_per_eval": B_JAC}
    audit = kernel_audit("linearize_tile")

    # ---- config 3: 65 536 x 100 nodes PER GPU, SoA outputs (7.7 GB written per launch: larger than L2)
    n_unique, rep = 655_360, 10
    x = torch.tensor(make_states(n_unique, seed=300 + rank, perturb_norm=1e-3).T, dtype=torch.float32, device=dev)
    u = torch.tensor(make_controls(n_unique, seed=300 + rank).T, dtype=torch.float32, device=dev)
    x, u = x.repeat(1, rep).contiguous(), u.repeat(1, rep).contiguous()
    n = x.shape[1]
    barrier()
    t_mark = clocks.mark()
    ms = rank_max(time_launches(torch, lambda: ac.linearize(x, u, DT, layout="soa"), K, 3, None))
    windows.append((t_mark, clocks.mark()))
    gbs = n * B_JAC / (ms * 1e-3) / 1e9
    ex = audit["executed_fp32_flops_per_unit"] if audit else None
    roof = {"kernel": "linearize_tile_kernel<64,soa> (64-node tiles, two blocks per SM)", "bound": "fp32 (reported in GB/s as north_star asks; intensity "
            "~25 flop/B is above the machine balance ~11)", "achieved": gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s",
            "frac": gbs / peaks["hbm_gbs"], "peak_source": peak_src, "bytes_per_eval": B_JAC,
            "traffic": audit["dram_bytes_per_unit"] * n if audit else None,
            "traffic_source": (f"profiles/{audit['capture']} via profiles/r2_kernel_audit.json, {audit['dram_bytes_per_unit']:.0f} B "
                               f"per eval scaled to this launch; CUDA sources "
                               f"{'CHANGED since' if audit['stale'] else 'unchanged since'} the capture") if audit else None}
    if ex:
        roof.update({"executed_flops_per_eval": ex, "tflops_executed": n * ex / (ms * 1e-3) / 1e12,
                     "frac_fp32_executed": n * ex / (ms * 1e-3) / 1e12 / fp32_peak_tflops,
                     "fp32_peak": fp32_peak_tflops})
    out["c3"] = {"workload": f"BASELINE config 3: 65 536 x 100 shooting nodes per GPU ({n * world} in total), A and B at every "
                             "node, SoA (7.7 GB written per GPU per launch: larger than L2)",
                 "scaling": "weak", "nodes_per_gpu": n, "ms_per_launch": ms, "value": n * world / (ms * 1e-3), "roofline": roof}
    out["value"] = out["c3"]["value"]
    del x, u

    # ---- config 5 EXACTLY: 8192 windows x 50 nodes in total, sharded over the ranks, AoS blocks [.,13,13], [.,13,7]
    from aircraft_b200.sharding import shard_range
    a5, b5 = shard_range(8192, rank, world)
    n5 = (b5 - a5) * 50
    xw = torch.tensor(make_states(n5, seed=500 + rank, perturb_norm=1e-3), dtype=torch.float32, device=dev).view(-1, 50, 13)
    uw = torch.tensor(make_controls(n5, seed=500 + rank), dtype=torch.float32, device=dev).view(-1, 50, 7)
    flush = torch.zeros(64 * 1024 * 1024, dtype=torch.float32, device=dev)
    barrier()
    ms5 = rank_max(time_launches(torch, lambda: ac.linearize(xw, uw, DT), K, 3, flush))
    out["c5"] = {"workload": "BASELINE config 5 exactly: 8192 windows x 50 nodes in total over the N GPUs "
                             f"({n5} nodes on this rank), state + control Jacobians, AoS",
                 "scaling": "strong", "nodes_total": 8192 * 50, "ms_per_launch": ms5, "value": 8192 * 50 / (ms5 * 1e-3),
                 "gbs": 8192 * 50 * B_JAC / (ms5 * 1e-3) / 1e9, "l2": "flushed between launches"}

    # ---- e2e: the same config-5 shard through the C-ABI host entry point with PINNED host buffers
    xh = torch.empty((n5, 13), dtype=torch.float32).pin_memory(); xh.copy_(xw.view(-1, 13))
    uh = torch.empty((n5, 7), dtype=torch.float32).pin_memory(); uh.copy_(uw.view(-1, 7))
    xn_h = torch.empty((n5, 13), dtype=torch.float32).pin_memory()
    A_h = torch.empty((n5, 13, 13), dtype=torch.float32).pin_memory()
    B_h = torch.empty((n5, 13, 7), dtype=torch.float32).pin_memory()
    lib, h = ac._lib, ac.handle

    def e2e_step():
        rc = lib.airc_linearize_host(h, xh.data_ptr(), uh.data_ptr(), DT, None, n5, 1, _cabi.FLAG_NORMALISE,
                                     _cabi.LAYOUT_AOS, xn_h.data_ptr(), A_h.data_ptr(), B_h.data_ptr())
        assert rc == 0
        return float(A_h[0, 3, 3])

    for _ in range(2):
        e2e_step()
    barrier()
    Ke = max(3, min(K, 10))
    t0 = time.perf_counter()
    for _ in range(Ke):
        e2e_step()
    el = rank_max(time.perf_counter() - t0)
    out["e2e"] = {"value": 8192 * 50 * Ke / el, "unit": "jacobian-evals/s", "workload": "config 5",
                  "h2d_bytes_per_step": n5 * 80, "d2h_bytes_per_step": n5 * 1092, "steps": Ke,
                  "path": "airc_linearize_host (C ABI; pinned host x, u in; x+, A, B out; PCIe-bound: 1 092 B per node)"}
    return out


def side_extras(torch, ac, dev, args):
    """Side lines measured on rank 0 only (device-timed, CUDA events)."""
    from aircraft_b200.workloads import make_controls, make_states
    out = {}
    rank = 0
    flush = torch.zeros(64 * 1024 * 1024, dtype=torch.float32, device=dev)   # 256 MB > 126 MB L2
    K = max(5, min(args.steps, 20))

    # config 4 shard with EXPLICIT control tensors (the form a caller with its own sequences uses)
    if args.workload == "c4":
        n4 = args.per_gpu or C4_TOTAL // C4_GPUS
        x1 = make_states(1, seed=4, level=True)
        x0 = torch.tensor(x1.T, dtype=torch.float32, device=dev).expand(13, n4).contiguous()
        U = device_control_sequences(torch, n4, H_ROLL, seed=100 + rank, device=dev)
        ms = time_launches(torch, lambda: ac.rollout(x0, U, DT, layout="soa", return_final=False, cost=True), K, 3, None)
        x0_h = torch.empty((13, n4), dtype=torch.float32).pin_memory(); x0_h.copy_(x0)
        U_h = torch.empty((H_ROLL, 7, n4), dtype=torch.float32).pin_memory(); U_h.copy_(U)
        x0_np, U_np = x0_h.numpy(), U_h.numpy()
        ac.rollout(x0_np, U_np, DT, layout="soa", return_final=False, cost=True)
        t0 = time.perf_counter()
        for _ in range(3):
            ac.rollout(x0_np, U_np, DT, layout="soa", return_final=False, cost=True)
        el = (time.perf_counter() - t0) / 3
        out["c4_explicit"] = {"workload": f"config 4 shard with explicit control tensors: {n4} x {H_ROLL}, U[H,7,B] = "
                                          f"{U_np.nbytes / 1e6:.0f} MB per launch (larger than L2)",
                              "value": n4 * H_ROLL / (ms * 1e-3), "unit": "state-steps/s", "ms_per_launch": ms,
                              "tflops": n4 * H_ROLL * F_STEP / (ms * 1e-3) / 1e12,
                              "e2e": {"value": n4 * H_ROLL / el, "unit": "state-steps/s",
                                      "h2d_bytes_per_step": x0_np.nbytes + U_np.nbytes, "d2h_bytes_per_step": 4 * n4,
                                      "path": "airc_rollout_host (PCIe-bound: the control tensor crosses the bus)"}}
        del x0, U, x0_h, U_h

    # config 1: ONE trajectory x 100 steps (the reference's own CPU-runnable case): latency per step
    x1 = torch.tensor(make_states(1, seed=1, level=True), dtype=torch.float32, device=dev)
    U1 = device_control_sequences(torch, 1, 100, seed=11, device=dev).permute(2, 0, 1).contiguous()   # [1,100,7]
    ms1 = time_launches(torch, lambda: ac.rollout(x1, U1, DT, return_trajectory=True), max(K, 20), 3, None)
    ms = time_launches(torch, lambda: ac.rollout(x1, U1, DT, return_trajectory=True, cooperative=True), max(K, 20), 3, None)
    out["c1"] = {"workload": "BASELINE config 1: 1 trajectory x 100 steps, full trajectory returned (latency case); latency "
                             "mode (8 lanes per trajectory, the default dispatch at this size; same launch with the "
                             "flag under ms_default)",
                 "ms_default": ms1,
                 "value": ms * 1e3 / 100, "unit": "us/state-step", "higher_is_better": False, "ms_per_launch": ms}
    if rank == 0:
        try:
            ref1 = _cpu_ref(1)
            xs, Us = make_states(1, seed=1, level=True).astype(np.float64), U1.double().cpu().numpy()
            ref1.rollout(xs, Us, DT)
            t0 = time.perf_counter()
            for _ in range(20):
                ref1.rollout(xs, Us, DT)
            out["c1"]["cpu_baseline"] = {"value": (time.perf_counter() - t0) / 20 / 100 * 1e6, "unit": "us/state-step",
                                         "cores": 1, "kind": "port",
                                         "sample": "20 repetitions of the same trajectory, oracle/cpu_ref.cpp float64"}
        except Exception as e:  # the checker is optional for this side line
            out["c1"]["cpu_baseline"] = {"unavailable": repr(e)}

    # The reference's only published timings (todo.md:38): IPOPT callbacks at 298 nodes with the
    # analytic DEFAULT coefficient model on one CPU thread.  The same numeric blocks on the GPU, same model.
    try:
        from aircraft_b200 import Aircraft
        from aircraft_b200.config import TRIM_COM
        acd = Aircraft.from_problem(com=TRIM_COM, device=dev.index, coeff_model_type="default",
                                    physical_integration_substeps=1, normalise=True)
        nn_ = 298
        xa = torch.tensor(make_states(nn_, seed=700 + rank), dtype=torch.float32, device=dev)
        ua = torch.tensor(make_controls(nn_, seed=700 + rank), dtype=torch.float32, device=dev)
        la = torch.randn(nn_, 13, device=dev, generator=torch.Generator(device=dev).manual_seed(700 + rank))
        xt = acd.state_update(xa.T.contiguous(), ua.T.contiguous(), DT).T.contiguous()
        Xw = torch.cat([xa[:1], xt], 0)[None]                    # a [1, 299, 13] guess, U [1, 298, 7]
        t_g = time_launches(torch, lambda: acd.state_update(xa.T, ua.T, DT), max(K, 20), 3, flush)
        t_jf = time_launches(torch, lambda: acd.defects(Xw, ua[None], DT), max(K, 20), 3, flush)
        # the same blocks through the C ABI on prepared buffers (the facade call above also slices / copies X into
        # x_k, x_{k+1}: three small torch kernels and their launch latency)
        from aircraft_b200 import _cabi as _cb
        rj, Aj, Bj = torch.empty_like(xa), torch.empty(nn_, 13, 13, device=dev), torch.empty(nn_, 13, 7, device=dev)
        stj = torch.cuda.current_stream().cuda_stream
        t_j = time_launches(torch, lambda: acd._lib.airc_defects(
            acd.handle, xa.data_ptr(), ua.data_ptr(), xt.data_ptr(), DT, None, nn_, 1, 1, _cb.LAYOUT_AOS, rj.data_ptr(),
            Aj.data_ptr(), Bj.data_ptr(), None, stj), max(K, 20), 3, flush)
        t_h = time_launches(torch, lambda: acd.lagrangian_hessian(xa, ua, la, DT), max(K, 20), 3, flush)
        out["mpc_blocks_default_298"] = {
            "workload": "298 shooting nodes, DEFAULT coefficient model, 1 substep: the dynamics blocks behind IPOPT's "
                        "nlp_g / nlp_jac_g / nlp_hess_l callbacks (values, defect Jacobian blocks A B, exact Hessian blocks)",
            "unit": "ms per evaluation", "higher_is_better": False,
            "state_update_ms": t_g, "defects_with_jacobian_blocks_ms": t_j, "defects_through_facade_ms": t_jf,
            "lagrangian_hessian_ms": t_h,
            "reference_published": {"nlp_g_ms": 0.65752, "nlp_jac_g_ms": 14.24, "nlp_hess_l_ms": 55.79,
                                    "source": "reference todo.md:38 via BASELINE.md section 2 (unknown CPU, 1 thread; whole-NLP "
                                              "callbacks incl. non-dynamics constraints: context, not a like-for-like ratio)"},
            "l2": "flushed between launches"}
    except Exception as e:   # side line only
        out["mpc_blocks_default_298"] = {"unavailable": repr(e)}

    # row f.1: exact Lagrangian-Hessian blocks at the reference's horizon (298 nodes, todo.md tables) and in bulk
    for name, nh in (("hessian_298", 298), ("hessian_bulk", 262_144)):
        x = torch.tensor(make_states(nh, seed=600 + rank, perturb_norm=1e-3).T, dtype=torch.float32, device=dev).contiguous()
        u = torch.tensor(make_controls(nh, seed=600 + rank).T, dtype=torch.float32, device=dev).contiguous()
        lam = torch.randn(13, nh, device=dev, generator=torch.Generator(device=dev).manual_seed(600 + rank))
        ms = time_launches(torch, lambda: ac.lagrangian_hessian(x, u, lam, DT, layout="soa", normalise=True), K, 3, flush)
        out[name] = {"workload": f"d2(lambda^T Phi)/d(x,u,dt)^2 [21x21] + gradient at {nh} shooting nodes (SoA)",
                     "kernel": "second-order jets, 64 threads per node (short horizon)"
                               if nh <= 4 * torch.cuda.get_device_properties(dev).multi_processor_count
                               else "forward over reverse, 8 lanes per node",
                     "value": nh / (ms * 1e-3), "unit": "hessian-evals/s", "ms_per_launch": ms,
                     "l2": "flushed between launches",
                     "reference_context": "nlp_hess_l 56-560 ms per IPOPT iteration at 298 nodes (reference todo.md tables, CPU)"}
        aud = kernel_audit("hessian2")
        if aud and name == "hessian_bulk":
            ex, by = aud["executed_fp32_flops_per_unit"], aud["dram_bytes_per_unit"]
            tf = nh * ex / (ms * 1e-3) / 1e12
            pk = fp32_peak(ac)
            out[name]["roofline"] = {
                "kernel": "hessian2_kernel<soa, 256> (forward over reverse on dual numbers, 8 lanes per node)", "bound": "fp32",
                "achieved": tf, "peak": pk, "unit": "TFLOP/s", "frac": tf / pk,
                "executed_flops_per_eval": ex, "dram_bytes_per_eval": by,
                "algorithmic_bytes_per_eval": 4 * (13 + 7 + 13 + 441 + 21),
                "issue_active_pct_under_ncu": aud.get("issue_active_pct"),
                "note": "latency bound rather than FMA bound: ~17 k straight-line instructions per thread with two warps "
                        "per scheduler at 255 registers, waiting on its own spills (profiles/r2_variants_hessian.txt); the "
                        "value parts are the same in the eight lanes of a node, so about a third of the executed flops "
                        "are replicated work",
                "source": f"profiles/{aud['capture']} via profiles/r2_kernel_audit.json (EXECUTED flops); CUDA sources "
                          f"{'CHANGED since' if aud['stale'] else 'unchanged since'} the capture"}
    return out


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_native(args)


if __name__ == "__main__":
    main()
