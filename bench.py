"""Benchmark of the validity-checking hot path (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W [--workload NAME] [--impl reference]

A step is one pass of the fused validity kernel (FK -> placements -> all enabled pairs) over
one batch of synthetic configurations per GPU; for N > 1 the step also all-gathers the
per-configuration flags over NCCL.  Prints ONE JSON line (rank 0).  ``--impl reference`` times
the CPU path (this repo's FP64 restatement of Pinocchio + coal: the reference's own
arithmetic lives in wheels that are not installable here) on the host cores instead.
"""

from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

METRIC = "collision-checked configs/sec (FK+all pairs)"
UNIT = "configs/s"
N_ROTATE = 8  # distinct input batches cycled through so no step re-reads an L2-resident batch


def measured_peaks() -> dict:
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        return {**json.loads(p.read_text()), "_source": "measured"}
    return {"hbm_gbs": 6650.0, "sm_max_mhz": 1965.0, "_source": "fallback"}


class ClockSampler:
    """SM clock and throttle reasons of one GPU, polled through NVML every millisecond by a
    thread while the timed region runs (a region of a few milliseconds is over before a single
    nvidia-smi process has answered).  Falls back to `nvidia-smi -lms` without pynvml."""

    REASONS = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.lines = index, None, []
        self.sm, self.mx, self.mask, self.nvml, self.handle = [], None, 0, None, None
        self._stop = threading.Event()
        self.thread = None
        try:
            import pynvml
            import torch

            pynvml.nvmlInit()
            try:
                uuid = "GPU-" + str(torch.cuda.get_device_properties(index).uuid)
                self.handle = pynvml.nvmlDeviceGetHandleByUUID(uuid.encode())
            except Exception:
                self.handle = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.mx = float(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
            self.nvml = pynvml
        except Exception:
            self.nvml = None

    def _poll(self):
        nv = self.nvml
        while not self._stop.is_set():
            try:
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(self.handle, nv.NVML_CLOCK_SM)))
                self.mask |= int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.handle))
            except Exception:
                try:
                    self.mask |= int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle))
                except Exception:
                    pass
            time.sleep(0.001)

    def start(self):
        if self.nvml is not None:
            self.thread = threading.Thread(target=self._poll, daemon=True)
            self.thread.start()
            return
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line)

    def stop(self) -> dict:
        if self.nvml is not None:
            self._stop.set()
            self.thread.join(timeout=1.0)
            reasons = sorted(k for k, bit in self.REASONS.items() if self.mask & bit)
            return {"sm_mhz": statistics.median(self.sm) if self.sm else None, "sm_max_mhz": self.mx,
                    "samples": len(self.sm), "reasons": reasons, "source": "nvml, 1 ms poll inside the timed region"}
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.lines:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons), "source": "nvidia-smi -lms 200"}


def host_threads() -> int:
    """All host cores this process may use (torchrun exports OMP_NUM_THREADS=1: ignore it)."""
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def reference_wheels() -> str:
    """BASELINE.md section 3 probe: are Pinocchio / coal / OMPL (the reference's own arithmetic)
    importable on this box?  They are not vendored; when absent the CPU arm is the oracle port."""
    import importlib.util

    found = [m for m in ("pinocchio", "hppfcl", "coal", "ompl") if importlib.util.find_spec(m) is not None]
    ref = os.path.isdir(os.path.join(os.path.dirname(os.path.abspath(__file__)), "baseline", "_ref"))
    return ("importable: " + ",".join(found) if found else "pinocchio/hppfcl/coal/ompl not importable") + (
        "; baseline/_ref present" if ref else "; no baseline/_ref") + " -> CPU arm = oracle port"


def cpu_baseline(desc, q_host, target_seconds: float = 12.0) -> dict:
    """The CPU oracle (boolean, stop at first collision, all host threads) on a bounded sample."""
    from oracle import oracle

    cores = host_threads()
    probe = q_host[: min(len(q_host), 1 << 14)]
    t0 = time.perf_counter()
    oracle.check_f32(desc, probe, cores)
    rate = len(probe) / (time.perf_counter() - t0)
    n = int(min(len(q_host), max(len(probe), rate * target_seconds)))
    t0 = time.perf_counter()
    oracle.check_f32(desc, q_host[:n], cores)
    dt = time.perf_counter() - t0
    return {"value": n / dt, "unit": UNIT, "cores": cores, "kind": "port", "reference_wheels": reference_wheels(),
            "sample": f"first {n} configurations of the step's batch, boolean stop-at-first, OpenMP {cores} threads, {dt:.1f} s"}


def run_reference(args, wl, rm, desc) -> None:
    """--impl reference: the CPU path on the host cores (rank 0 only)."""
    import numpy as np

    from oracle import oracle
    from ramp_b200 import workloads

    if int(os.environ.get("RANK", "0")) != 0:
        return
    cores = host_threads()
    q = workloads.sample_q(rm.model, 1 << 18, seed=0xB200)
    t0 = time.perf_counter()
    oracle.check_f32(desc, q[: 1 << 13], cores)
    rate = (1 << 13) / (time.perf_counter() - t0)
    total = args.steps + args.warmup
    n = int(min(len(q), max(1 << 12, rate * min(20.0, 150.0 / max(total, 1)))))
    for _ in range(args.warmup):
        oracle.check_f32(desc, q[:n], cores)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        oracle.check_f32(desc, q[:n], cores)
    dt = time.perf_counter() - t0
    value = n * args.steps / dt
    sample = f"{n} of the workload's configurations per step (bounded sample), boolean stop-at-first, OpenMP {cores} threads"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": wl.name, "robot": wl.robot, "scene": wl.scene, "note": wl.note,
                   "model": "stand-in" if "stand-in" in wl.note else "in-tree"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample,
                         "reference_wheels": reference_wheels()},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "reference_note": "Pinocchio 3.7.0 / coal 3.0.1 are not installable offline; this is the repo's FP64 "
                          "restatement of that path (oracle/ramp_oracle.c)",
    }))


def run_fk(args, wl, rm, dev) -> None:
    """FK with poses stored (ramp_fk_batch): the HBM-bound kernel of the path."""
    import torch

    from ramp_b200 import engine, workloads

    n = args.n or min(wl.n, 1 << 22)
    ld = workloads.padded_nq(rm.model)
    nj = rm.model.njoints - 1
    gpu = rm.gpu
    q = workloads.sample_q_torch(rm.model, n, seed=0xB200, device=dev)
    out = torch.empty((n, nj, 3, 4), dtype=torch.float32, device=dev)
    lib, h, st = gpu.lib, gpu.handle, torch.cuda.current_stream(dev).cuda_stream
    for _ in range(args.warmup):
        lib.ramp_fk_batch(h, q.data_ptr(), n, ld, 0, out.data_ptr(), st)
    torch.cuda.synchronize()
    l0 = engine.launch_count()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(args.steps):
        lib.ramp_fk_batch(h, q.data_ptr(), n, ld, 0, out.data_ptr(), st)
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / args.steps
    peaks = measured_peaks()
    bytes_per_cfg = 4 * ld + 48 * nj
    gbs = bytes_per_cfg * n / (ms * 1e-3) / 1e9
    print(json.dumps({
        "metric": "FK configs/sec (poses stored)", "value": n / (ms * 1e-3), "unit": UNIT, "n_gpus": 1, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": wl.name + ":fk", "robot": wl.robot, "n": n, "joints": nj, "ld_q": ld,
                   "l2": f"output {48 * nj * n / 2**20:.0f} MiB per step streams through L2 (> 126 MB)"},
        "roofline": {"bound": "hbm", "achieved": gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": gbs / peaks["hbm_gbs"],
                     "traffic": None, "kernel": "fk_kernel<1>", "algorithmic_bytes_per_config": bytes_per_cfg,
                     "peak_source": peaks["_source"]},
        "gpu_launches": engine.launch_count() - l0,
    }))


def run_edges(args, wl, rm, desc, dev) -> None:
    """BASELINE config 3 (planner edges through the batched motion validator), single GPU:
    a step is one ramp_check_edges call over E random edges; value = interpolated states
    checked per second.  Also times one RRTConnect query through the same validator."""
    import numpy as np
    import torch

    from ramp_b200 import MotionPlanner, RobotState, engine, spaces

    planner = MotionPlanner(rm, "arm")
    start = RobotState.from_named_state(rm, "arm", "home")
    planner.set_reference_state(start)
    E = args.n or wl.n
    rng = np.random.default_rng(0xB200)
    q0 = spaces.sample_uniform(planner.space, E, rng)
    lo = np.array([s["lo"][0] for s in planner.space.subspaces])
    hi = np.array([s["hi"][0] for s in planner.space.subspaces])
    q1 = np.clip(q0 + rng.normal(scale=0.4, size=q0.shape), lo, hi)
    d0, d1 = torch.from_numpy(q0.astype(np.float32)).to(dev), torch.from_numpy(q1.astype(np.float32)).to(dev)
    gpu = rm.gpu
    res = gpu.check_edges(d0, d1, planner.space)
    states = int(res["nd"].clamp(min=1).sum().item())
    for _ in range(args.warmup):
        gpu.check_edges(d0, d1, planner.space)
    torch.cuda.synchronize()
    l0 = engine.launch_count()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(args.steps):
        res = gpu.check_edges(d0, d1, planner.space)
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / args.steps
    launches = engine.launch_count() - l0
    # end to end: host end points in, host verdicts out (ramp_check_edges_host)
    h0, h1 = q0.astype(np.float32), q1.astype(np.float32)
    gpu.check_edges_host(h0, h1, planner.space)
    e2e_steps = max(3, min(args.steps, 10))
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        gpu.check_edges_host(h0, h1, planner.space)
    e2e_value = states * e2e_steps / (time.perf_counter() - t0)
    t0 = time.perf_counter()
    path = planner.plan(start, [0.0, 0.25, 0.0, -1.25, 0.0, 1.5, 0.0], timeout=5.0)
    plan_s = time.perf_counter() - t0
    print(json.dumps({
        "metric": METRIC, "value": states / (ms * 1e-3), "unit": UNIT, "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": wl.name, "robot": wl.robot, "scene": wl.scene, "note": wl.note, "model": "stand-in",
                   "edges": E, "states": states, "mean_nd": states / E, "edges_per_s": E / (ms * 1e-3),
                   "valid_edge_fraction": float(res["valid"].float().mean().item())},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h0.nbytes + h1.nbytes), "d2h_bytes_per_step": 9 * E,
                "api": "ramp_check_edges_host (C ABI, host buffers)", "steps": e2e_steps},
        "gpu_launches": launches,
        "planning": {"planner": "built-in RRTConnect over GpuMotionValidator", "found": path is not None,
                     "seconds": plan_s, "waypoints": 0 if path is None else len(path),
                     "edges_checked": planner.validator.valid_count + planner.validator.invalid_count,
                     "states_checked": planner.validator.states_checked},
    }))


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="ur5e_1m")
    ap.add_argument("--n", "--rows", dest="n", type=int, default=0, help="override configurations per step per GPU (--rows under torchrun, whose own parser claims --n)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--kernel", default="check", choices=["check", "fk"], help="fk: time ramp_fk_batch (poses stored)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    import __graft_entry__

    rank = int(os.environ.get("RANK", "0"))
    if rank == 0:
        __graft_entry__.build()

    from ramp_b200 import workloads
    from ramp_b200.desc import ModelDesc

    wl = workloads.WORKLOADS[args.workload]
    rm, _ = workloads.build(args.workload)
    desc = ModelDesc(rm.model, rm.collision_model)
    if args.impl == "reference":
        run_reference(args, wl, rm, desc)
        return

    import numpy as np
    import torch
    import torch.distributed as dist

    from ramp_b200 import engine

    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU fallback); use --impl reference for the CPU path")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
        dist.barrier()

    if args.workload == "fr3_edges":
        run_edges(args, wl, rm, desc, dev)
        return
    if args.kernel == "fk":
        run_fk(args, wl, rm, dev)
        return
    n = args.n or wl.n
    gpu = rm.gpu
    ld = workloads.padded_nq(rm.model)
    # N_ROTATE distinct batches: 8 x n x ld x 4 B is well above the 126 MB L2 for n >= 2^20
    batches = [workloads.sample_q_torch(rm.model, n, seed=0xB200 + 97 * rank + 13 * i, device=dev) for i in range(N_ROTATE)]
    # Two result buffers: the NCCL all-gather of step i's flags runs on NCCL's stream while the
    # kernels of step i+1 fill the other buffer; step i+2 waits for it before reusing the buffer.
    hits = [torch.empty(n, dtype=torch.uint8, device=dev) for _ in range(2)]
    hit = hits[0]
    gathered = [torch.empty(n * world, dtype=torch.uint8, device=dev) for _ in range(2)] if world > 1 else None
    pending = [None, None]

    def step(i: int):
        b = i & 1
        if pending[b] is not None:
            pending[b].wait()  # stream-side wait: the gather that read hits[b] is done
            pending[b] = None
        gpu.check_raw(batches[i % N_ROTATE], hits[b])
        if world > 1:
            pending[b] = dist.all_gather_into_tensor(gathered[b], hits[b], async_op=True)

    def drain():
        for b in range(2):
            if pending[b] is not None:
                pending[b].wait()
                pending[b] = None

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for i in range(args.warmup):
        step(i)
    drain()
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    launches0 = engine.launch_count()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    barrier()
    ev[0].record()
    for i in range(args.steps):
        step(i)
        if i + 1 == args.steps:
            drain()  # every step's gather has completed before the last event
        ev[i + 1].record()
    barrier()
    launches = engine.launch_count() - launches0
    clocks = sampler.stop()
    total_ms = ev[0].elapsed_time(ev[-1])
    t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    colliding = float(hit.float().mean().item())

    # kernel-only duration: the same launches again, one event pair per launch, no collective
    kt = []
    for i in range(min(args.steps, 10)):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        gpu.check_raw(batches[i % N_ROTATE], hit)
        b.record()
        b.synchronize()
        kt.append(a.elapsed_time(b))
    kernel_ms = statistics.mean(kt)

    # end to end: the C-ABI host-buffer call (ramp_check_host): H2D of q, kernel, D2H of flags.
    # The step's inputs start in page-locked host memory as plain [n, nq] rows (no padding).
    from ramp_b200.engine import pinned_empty
    nq = rm.model.nq
    q_host = pinned_empty((n, nq), np.float32)
    q_host[:] = batches[0][:, :nq].cpu().numpy()
    hit_host = pinned_empty((n,), np.uint8)
    e2e_steps = max(3, min(args.steps, 10))
    gpu.check_host_into(q_host, hit_host)
    assert np.array_equal(hit_host.astype(bool), gpu.check(batches[0])["hit"].cpu().numpy()), "e2e flags differ"
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        gpu.check_host_into(q_host, hit_host)
    barrier()
    e2e_s = time.perf_counter() - t0
    te = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_value = n * world * e2e_steps / float(te.item())

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peaks = measured_peaks()
    from oracle import oracle

    sample = q_host[:4096, : rm.model.nq].astype(np.float64)
    flops_per_cfg = float(oracle.flops(desc, sample).mean())
    bytes_per_cfg = 4 * ld + 1
    sm_count = torch.cuda.get_device_properties(dev).multi_processor_count
    fp32_peak = sm_count * 128 * 2 * peaks["sm_max_mhz"] * 1e6 / 1e12
    achieved_tf = flops_per_cfg * n / (kernel_ms * 1e-3) / 1e12
    achieved_gbs = bytes_per_cfg * n / (kernel_ms * 1e-3) / 1e9
    traffic = None
    tp = ROOT / "profiles" / "traffic.json"
    if tp.exists():
        traffic = json.loads(tp.read_text()).get(args.workload)
    ref = oracle.check(desc, sample[:2048])
    got = gpu.check(batches[0][:2048])["hit"].cpu().numpy()
    outside = np.abs(ref["min_dist"]) > 1e-4
    out = {
        "metric": METRIC, "value": n * world * args.steps / (total_ms * 1e-3), "unit": UNIT, "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": wl.name, "robot": wl.robot, "scene": wl.scene, "note": wl.note,
                   "model": "stand-in" if "stand-in" in wl.note else "in-tree",
                   "n_per_gpu": n, "pairs": desc.n_pairs, "geoms": desc.n_geoms, "ld_q": ld,
                   "colliding_fraction": colliding,
                   "l2": f"{N_ROTATE} rotating input batches of {n * ld * 4 / 2**20:.0f} MiB each (inputs > 126 MB L2)",
                   "parallelism": f"config shards x{world}, geometry replicated, flags all-gathered per step (NCCL, overlapped with the next step)"
                                  + (", NCCL all_gather of flags per step" if world > 1 else "")},
        "roofline": {"bound": "fp32", "achieved": achieved_tf, "peak": fp32_peak, "unit": "TFLOP/s",
                     "frac": achieved_tf / fp32_peak, "traffic": traffic,
                     "kernel": "check_kernel + gjk_queue_kernel (one ramp_check_batch call; check_kernel is ~2/3 of it)",
                     "kernel_ms": kernel_ms, "algorithmic_flops_per_config": flops_per_cfg,
                     "peak_source": f"derived: {sm_count} SMs x 128 FP32 lanes x 2 x sm_max_mhz "
                                    f"({peaks['_source']} MEASURED_PEAKS.json); no measured FP32 peak exists",
                     "hbm": {"achieved": achieved_gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                             "frac": achieved_gbs / peaks["hbm_gbs"], "algorithmic_bytes_per_config": bytes_per_cfg,
                             "peak_source": peaks["_source"]}},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": n * nq * 4, "d2h_bytes_per_step": n,
                "api": "ramp_check_host (C ABI, page-locked host buffers, rows of nq floats)", "steps": e2e_steps},
        "gpu_launches": launches,
        "clocks": clocks,
        "parity": {"checked": int(len(got)), "in_band": int((~outside).sum()),
                   "mismatch_outside_band": int((got != ref["hit"])[outside].sum())},
    }
    if world == 1 and not args.no_cpu_baseline:
        out["cpu_baseline"] = cpu_baseline(desc, q_host)
    print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
