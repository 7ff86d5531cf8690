"""Benchmark of the validity-checking hot path (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W [--impl reference] [--workload NAME] [--rows n] [--no-subs]

Headline (default ``panda_64m``, BASELINE.json configs[4]: the robot ``north_star``'s target is
quoted on): a STEP is ``passes`` back-to-back calls of the product API on this GPU's shard --
``ramp_b200.distributed.ShardedChecker.submit``: ramp_check_batch (FK -> placements -> every
enabled pair) over a batch of synthetic configurations, the flags packed to bits and, for N > 1,
all-gathered over NCCL while the next batch is being checked.  ``value`` = configurations all
ranks checked / device time (CUDA events, max over ranks).  ``e2e`` is the same workload through
the C-ABI host-buffer call (ramp_check_host: H2D of q, kernels, D2H of flags inside the timed
region).  At N = 1 the line also carries ``workloads``: BASELINE configs 1, 2, 4, the same Panda
with non-capsule hulls, FK with poses stored, and config 3 (planner edges), each with its own
roofline / e2e / parity.

``--impl reference`` times the CPU path on the host cores instead (this repo's FP64 restatement
of Pinocchio + coal: the reference's own arithmetic lives in wheels that are not installable
here) and never loads the CUDA library.  Prints ONE JSON line (rank 0).
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
HEADLINE = "panda_64m"
L2_BYTES = 126 << 20
MIN_STEP_MS = 25.0  # a step is at least this much GPU work (enough clock samples under load)

# (workload, rows per pass per GPU, note) measured next to the headline at N = 1
SUBS = [("panda_4096", 4096), ("ur5e_1m", 1 << 20), ("h1_16m", 1 << 24), ("panda_boxy_64m", 1 << 25)]


def measured_peaks() -> dict:
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        return {**json.loads(p.read_text()), "_source": "measured"}
    return {"hbm_gbs": 6650.0, "sm_max_mhz": 1965.0, "_source": "fallback"}


class ClockSampler:
    """SM clock and throttle reasons of one GPU, polled through NVML every millisecond by a
    thread while the timed region runs.  Falls back to `nvidia-smi -lms` without pynvml."""

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
    ref = (ROOT / "baseline" / "_ref").is_dir()
    return ("importable: " + ",".join(found) if found else "pinocchio/hppfcl/coal/ompl not importable") + (
        "; baseline/_ref present" if ref else "; no baseline/_ref") + " -> CPU arm = oracle port"


# ------------------------------------------------------------------ workload description
def plan_passes(name: str, n: int, passes_arg: int) -> int:
    """Calls per step so that a step is >= MIN_STEP_MS of GPU work (rates of round 1, rounded up)."""
    if passes_arg > 0:
        return passes_arg
    per_s = {"panda": 3.5e9, "ur5e": 2.3e9, "h1": 0.9e9, "rrr": 3e9, "fr3": 3e9}.get(name.split("_")[0], 2e9)
    if n <= 8192:
        per_s = n / 40e-6  # latency regime: one call is a launch, ~40 us
    return max(1, int(-(-MIN_STEP_MS * 1e-3 * per_s // n)))


def n_input_batches(batch_bytes: int) -> int:
    return 2 if batch_bytes > 2 * L2_BYTES else min(16, max(2, -(-(3 * L2_BYTES) // batch_bytes)))


def l2_note(batch_bytes: int) -> str:
    nb = n_input_batches(batch_bytes)
    if batch_bytes > 2 * L2_BYTES:
        return f"every pass reads a {batch_bytes / 2**20:.0f} MiB batch of its own ({nb} alternate): inputs larger than the 126 MB L2"
    if nb * batch_bytes > 2 * L2_BYTES:
        return f"passes rotate through {nb} distinct batches of {batch_bytes / 2**20:.0f} MiB ({nb * batch_bytes / 2**20:.0f} MiB > 2x the 126 MB L2)"
    return (f"latency regime: {nb} distinct batches of {batch_bytes / 2**10:.0f} KiB stay L2-resident "
            "(a call is launch-bound, not bandwidth-bound; the inputs cannot exceed L2 at this batch size)")


def config_for(wl, desc, model, n: int, passes: int, world: int) -> dict:
    """The workload's description: identical in the GPU arm and in the reference arm."""
    ld = (model.nq + 3) & ~3
    batch_bytes = n * ld * 4
    return {
        "workload": wl.name, "robot": wl.robot, "scene": wl.scene, "note": wl.note,
        "model": "stand-in" if "stand-in" in wl.note else "in-tree",
        "n_per_gpu": n, "passes_per_step": passes, "rows_per_step_per_gpu": n * passes,
        "pairs": desc.n_pairs, "geoms": desc.n_geoms, "nq": model.nq, "ld_q": ld,
        "l2": l2_note(batch_bytes),
        "parallelism": f"config shards x{world}, geometry replicated, flags gathered as bits per pass (NCCL all_gather, overlapped with the next pass)",
    }


def cpu_leg(desc, q_host, target_seconds: float) -> dict:
    """The CPU oracle (boolean, stop at first collision, all host threads) on a bounded sample."""
    from oracle import oracle

    cores = host_threads()
    probe = q_host[: min(len(q_host), 1 << 13)]
    t0 = time.perf_counter()
    oracle.check_f32(desc, probe, cores)
    rate = len(probe) / (time.perf_counter() - t0)
    n = int(min(len(q_host), max(len(probe), rate * target_seconds)))
    t0 = time.perf_counter()
    oracle.check_f32(desc, q_host[:n], cores)
    dt = time.perf_counter() - t0
    return {"value": n / dt, "unit": UNIT, "cores": cores, "kind": "port", "reference_wheels": reference_wheels(),
            "sample": f"first {n} configurations of the workload's batch, boolean stop-at-first, OpenMP {cores} threads, {dt:.1f} s"}


def run_reference(args) -> None:
    """--impl reference: the CPU path on the host cores (rank 0 only).  Loads the oracle and the
    pure-Python model description only -- never libramp_b200.so."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    from oracle import oracle
    from ramp_b200 import workloads
    from ramp_b200.desc import ModelDesc

    oracle.build()
    name = args.workload or HEADLINE
    wl = workloads.WORKLOADS[name]
    rm, _ = workloads.build(name)
    desc = ModelDesc(rm.model, rm.collision_model)
    n = args.n or wl.n
    passes = plan_passes(name, n, args.passes)
    cores = host_threads()
    q = workloads.sample_q(rm.model, 1 << 18, seed=0xB200)
    t0 = time.perf_counter()
    oracle.check_f32(desc, q[: 1 << 13], cores)
    rate = (1 << 13) / (time.perf_counter() - t0)
    total = args.steps + args.warmup
    sample = int(min(len(q), max(1 << 12, rate * min(20.0, 150.0 / max(total, 1)))))
    for _ in range(args.warmup):
        oracle.check_f32(desc, q[:sample], cores)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        oracle.check_f32(desc, q[:sample], cores)
    dt = time.perf_counter() - t0
    value = sample * args.steps / dt
    note = (f"each step times {sample} of the step's {n * passes} configurations (bounded sample of the same workload), "
            f"boolean stop-at-first, OpenMP {cores} threads")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": config_for(wl, desc, rm.model, n, passes, args.gpus),
        "sample_rows_per_step": sample,
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": note,
                         "reference_wheels": reference_wheels()},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "reference_note": "Pinocchio 3.7.0 / coal 3.0.1 are not installable offline; this is the repo's FP64 "
                          "restatement of that path (oracle/ramp_oracle.c)",
    }))


# ------------------------------------------------------------------ GPU arm
def profile_constants(name: str, n: int) -> tuple[dict | None, dict | None]:
    """DRAM traffic and hardware counters of the call's kernels from this round's ncu captures
    (profiles/traffic.json, profiles/hw_counters.json): constants of a profile, not measured in
    this run; traffic is scaled to this run's rows per launch."""
    traffic = hw = None
    tp, hp = ROOT / "profiles" / "traffic.json", ROOT / "profiles" / "hw_counters.json"
    if tp.exists():
        t = json.loads(tp.read_text()).get(name)
        if isinstance(t, dict) and t.get("rows"):
            k = n / t["rows"]
            traffic = {"bytes_per_call": k * (t["check_kernel_bytes"] + t["gjk_queue_kernel_bytes"]),
                       "check_kernel": k * t["check_kernel_bytes"], "gjk_queue_kernel": k * t["gjk_queue_kernel_bytes"],
                       "source": f"profile constant ({t.get('source', 'ncu --set full')}), scaled from {t['rows']} rows"}
    if hp.exists():
        hw = json.loads(hp.read_text()).get(name)
    return traffic, hw


def measure_check(name: str, n: int, passes: int, steps: int, warmup: int, *, world: int, rank: int, dev,
                  cpu_seconds: float, e2e_steps: int) -> dict:
    """One workload through the product API; returns the JSON fields of its line."""
    import numpy as np
    import torch
    import torch.distributed as dist

    from ramp_b200 import engine, workloads
    from ramp_b200.desc import ModelDesc
    from ramp_b200.distributed import ShardedChecker
    from ramp_b200.engine import pinned_empty

    wl = workloads.WORKLOADS[name]
    rm, _ = workloads.build(name)
    desc = ModelDesc(rm.model, rm.collision_model)
    gpu = rm.gpu
    ld = workloads.padded_nq(rm.model)
    batch_bytes = n * ld * 4
    n_batches = n_input_batches(batch_bytes)
    batches = [workloads.sample_q_torch(rm.model, n, seed=0xB200 + 97 * rank + 13 * i, device=dev) for i in range(n_batches)]
    sc = ShardedChecker(gpu, n)
    last = [0]

    def step(i: int):
        for p in range(passes):
            last[0] = sc.submit(batches[(i * passes + p) % n_batches])

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for i in range(warmup):
        step(i)
        if i == 0:
            torch.cuda.synchronize()  # the library sizes its work queue from the first large call's item counts
    sc.drain()
    barrier()
    gpu.counters(reset=True)
    sampler = ClockSampler(dev.index or 0)
    sampler.start()
    launches0 = engine.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record()
    for i in range(steps):
        step(i)
    sc.drain()  # every pass's gather has completed before the closing event
    ev1.record()
    barrier()
    launches = engine.launch_count() - launches0
    clocks = sampler.stop()
    total_ms = ev0.elapsed_time(ev1)
    t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    counters = gpu.counters()
    rows_timed = n * passes * steps
    # the gathered bits of the last pass against this rank's own flags: the exchange is read, not discarded
    local = sc.local_flags(last[0]).bool()
    gathered = sc.result(last[0])
    gather_ok = bool(torch.equal(gathered[rank * n:(rank + 1) * n], local)) and gathered.shape[0] == n * world
    colliding = float(local.float().mean().item())

    # end to end: the C-ABI host-buffer call (ramp_check_host): H2D of q, kernels, D2H of flags.
    # The inputs start in page-locked host memory as plain [n, nq] rows (no padding).
    nq = rm.model.nq
    q_host = pinned_empty((n, nq), np.float32)
    torch.from_numpy(q_host).copy_(batches[0][:, :nq])
    hit_host = pinned_empty((n,), np.uint8)
    gpu.check_host_into(q_host, hit_host)
    e2e_same = bool(np.array_equal(hit_host.astype(bool), gpu.check(batches[0])["hit"].cpu().numpy()))
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        gpu.check_host_into(q_host, hit_host)
    barrier()
    te = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_value = n * world * e2e_steps / float(te.item())

    out = {"workload": name, "metric": METRIC, "unit": UNIT, "value": rows_timed * world / (total_ms * 1e-3),
           "ms_per_step": total_ms / steps, "steps": steps, "warmup": warmup,
           "config": config_for(wl, desc, rm.model, n, passes, world),
           "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": n * nq * 4, "d2h_bytes_per_step": n,
                   "api": "ramp_check_host (C ABI, page-locked host buffers, rows of nq floats); an e2e step is one call over n_per_gpu rows",
                   "steps": e2e_steps, "flags_equal_device_call": e2e_same},
           "gpu_launches": launches, "clocks": clocks,
           "measured": {"colliding_fraction": colliding, "gjk_items_per_config": counters["gjk_items"] / max(rows_timed, 1),
                        "gather_checked": gather_ok, "input_batches": n_batches}}
    if rank != 0:
        return out

    from oracle import oracle

    peaks = measured_peaks()
    sample = q_host[:4096, :nq].astype(np.float64)
    flops_stop = float(oracle.flops(desc, sample, stop_at_first=True).mean())
    flops_all = float(oracle.flops(desc, sample).mean())
    call_ms = total_ms / (steps * passes)  # one ramp_check_batch call (+ the flag packing), back to back on one stream
    sm_count = torch.cuda.get_device_properties(dev).multi_processor_count
    fp32_peak = sm_count * 128 * 2 * peaks["sm_max_mhz"] * 1e6 / 1e12
    achieved_tf = flops_stop * n / (call_ms * 1e-3) / 1e12
    bytes_per_cfg = 4 * ld + 1
    achieved_gbs = bytes_per_cfg * n / (call_ms * 1e-3) / 1e9
    traffic, hw = profile_constants(name, n)
    out["roofline"] = {
        "bound": "fp32", "achieved": achieved_tf, "peak": fp32_peak, "unit": "TFLOP/s", "frac": achieved_tf / fp32_peak,
        "traffic": None if traffic is None else traffic["bytes_per_call"], "traffic_detail": traffic,
        "kernel": "check_kernel + gjk_queue_kernel (one ramp_check_batch call)",
        "kernel_ms": call_ms, "kernel_ms_how": "CUDA events around the timed region on the launching stream / calls in it",
        "algorithmic_flops_per_config": flops_stop,
        "numerator": "SURVEY 8(d) formula on the oracle's GJK iteration counts, pair list of a configuration cut at its "
                     "first colliding pair (what the timed boolean call and the reference compute)",
        "algorithmic_flops_per_config_all_pairs": flops_all,
        "hw_counters_profile_constant": hw,
        "peak_source": f"derived: {sm_count} SMs x 128 FP32 lanes x 2 x sm_max_mhz ({peaks['_source']} MEASURED_PEAKS.json); "
                       "no measured FP32 peak exists",
        "hbm": {"achieved": achieved_gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": achieved_gbs / peaks["hbm_gbs"],
                "algorithmic_bytes_per_config": bytes_per_cfg, "peak_source": peaks["_source"]}}
    ref = oracle.check(desc, sample[:2048])
    got = gpu.check(batches[0][:2048])["hit"].cpu().numpy()
    outside = np.abs(ref["min_dist"]) > 1e-4
    out["parity"] = {"checked": int(len(got)), "in_band": int((~outside).sum()),
                     "mismatch_outside_band": int((got != ref["hit"])[outside].sum())}
    if cpu_seconds > 0:
        out["cpu_baseline"] = cpu_leg(desc, q_host, cpu_seconds)
    return out


def measure_fk(name: str, n: int, steps: int, warmup: int, dev) -> dict:
    """FK with poses stored (ramp_fk_batch): the HBM-bound kernel of the path."""
    import numpy as np
    import torch

    from oracle import oracle
    from ramp_b200 import engine, workloads
    from ramp_b200.desc import ModelDesc

    wl = workloads.WORKLOADS[name]
    rm, _ = workloads.build(name)
    ld = workloads.padded_nq(rm.model)
    nj = rm.model.njoints - 1
    gpu = rm.gpu
    q = workloads.sample_q_torch(rm.model, n, seed=0xB200, device=dev)
    out = torch.empty((n, nj, 3, 4), dtype=torch.float32, device=dev)
    lib, h, st = gpu.lib, gpu.handle, torch.cuda.current_stream(dev).cuda_stream
    for _ in range(warmup):
        lib.ramp_fk_batch(h, q.data_ptr(), n, ld, 0, out.data_ptr(), st)
    torch.cuda.synchronize()
    l0 = engine.launch_count()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        lib.ramp_fk_batch(h, q.data_ptr(), n, ld, 0, out.data_ptr(), st)
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / steps
    peaks = measured_peaks()
    bytes_per_cfg = 4 * ld + 48 * nj
    gbs = bytes_per_cfg * n / (ms * 1e-3) / 1e9
    ref = oracle.fk(ModelDesc(rm.model, rm.collision_model), q[:2048, : rm.model.nq].cpu().numpy().astype(np.float64))
    err = float(np.abs(out[:2048].cpu().numpy() - ref).max())
    return {"workload": name + ":fk", "metric": "FK configs/sec (poses stored)", "unit": UNIT, "value": n / (ms * 1e-3),
            "ms_per_step": ms, "steps": steps, "warmup": warmup,
            "config": {"workload": name + ":fk", "robot": wl.robot, "n": n, "joints": nj, "ld_q": ld,
                       "l2": f"output {48 * nj * n / 2**20:.0f} MiB per step streams through L2 (> 126 MB)"},
            "roofline": {"bound": "hbm", "achieved": gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": gbs / peaks["hbm_gbs"],
                         "traffic": None, "kernel": "fk_kernel<1>", "algorithmic_bytes_per_config": bytes_per_cfg,
                         "peak_source": peaks["_source"]},
            "parity": {"checked": 2048, "max_pose_error": err, "tolerance": 1e-5},
            "gpu_launches": engine.launch_count() - l0}


def measure_edges(name: str, n_edges: int, steps: int, warmup: int, dev, *, seeds: int, cpu_queries: int) -> dict:
    """BASELINE config 3 (planner edges through the batched motion validator), single GPU: a step
    is one ramp_check_edges call over E random edges; value = interpolated states checked per
    second.  Then ``seeds`` RRTConnect queries through the same validator (edge by edge, the way
    OMPL calls checkMotion) next to the same queries through a per-state CPU validator."""
    import numpy as np
    import torch

    from ramp_b200 import MotionPlanner, RobotState, engine, spaces, workloads

    wl = workloads.WORKLOADS[name]
    rm, _ = workloads.build(name)
    planner = MotionPlanner(rm, "arm")
    start = RobotState.from_named_state(rm, "arm", "home")
    planner.set_reference_state(start)
    rng = np.random.default_rng(0xB200)
    q0 = spaces.sample_uniform(planner.space, n_edges, rng)
    lo = np.array([s["lo"][0] for s in planner.space.subspaces])
    hi = np.array([s["hi"][0] for s in planner.space.subspaces])
    q1 = np.clip(q0 + rng.normal(scale=0.4, size=q0.shape), lo, hi)
    d0, d1 = torch.from_numpy(q0.astype(np.float32)).to(dev), torch.from_numpy(q1.astype(np.float32)).to(dev)
    gpu = rm.gpu
    res = gpu.check_edges(d0, d1, planner.space)
    states = int(res["nd"].clamp(min=1).sum().item())
    for _ in range(warmup):
        gpu.check_edges(d0, d1, planner.space)
    torch.cuda.synchronize()
    l0 = engine.launch_count()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        res = gpu.check_edges(d0, d1, planner.space)
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / steps
    launches = engine.launch_count() - l0
    h0, h1 = q0.astype(np.float32), q1.astype(np.float32)
    gpu.check_edges_host(h0, h1, planner.space)
    e2e_steps = max(3, min(steps, 10))
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        gpu.check_edges_host(h0, h1, planner.space)
    e2e_value = states * e2e_steps / (time.perf_counter() - t0)
    out = {"workload": name, "metric": "interpolated states checked/sec (planner edges)", "unit": "states/s",
           "value": states / (ms * 1e-3), "ms_per_step": ms, "steps": steps, "warmup": warmup,
           "config": {"workload": wl.name, "robot": wl.robot, "scene": wl.scene, "note": wl.note, "model": "stand-in",
                      "edges": n_edges, "states": states, "mean_nd": states / n_edges, "edges_per_s": n_edges / (ms * 1e-3),
                      "valid_edge_fraction": float(res["valid"].float().mean().item())},
           "e2e": {"value": e2e_value, "unit": "states/s", "h2d_bytes_per_step": int(h0.nbytes + h1.nbytes),
                   "d2h_bytes_per_step": 9 * n_edges, "api": "ramp_check_edges_host (C ABI, host buffers)", "steps": e2e_steps},
           "gpu_launches": launches}
    if n_edges < (1 << 18):
        # the same call on a roadmap-sized batch (2^18 edges, one slice): the throughput regime
        big = 1 << 18
        b0 = spaces.sample_uniform(planner.space, big, rng)
        b1 = np.clip(b0 + rng.normal(scale=0.4, size=b0.shape), lo, hi)
        e0, e1 = torch.from_numpy(b0.astype(np.float32)).to(dev), torch.from_numpy(b1.astype(np.float32)).to(dev)
        res = gpu.check_edges(e0, e1, planner.space)
        big_states = int(res["nd"].clamp(min=1).sum().item())
        for _ in range(2):
            gpu.check_edges(e0, e1, planner.space)
        torch.cuda.synchronize()
        a.record()
        for _ in range(steps):
            gpu.check_edges(e0, e1, planner.space)
        b.record()
        torch.cuda.synchronize()
        big_ms = a.elapsed_time(b) / steps
        out["large_batch"] = {"edges": big, "states": big_states, "ms_per_step": big_ms, "states_per_s": big_states / (big_ms * 1e-3),
                              "edges_per_s": big / (big_ms * 1e-3)}
    if seeds > 0:
        out["planning"] = planning_leg(seeds, cpu_queries)
    return out


class CpuStateValidator:
    """The reference's arrangement for motion validity, on the CPU oracle: OMPL's discrete motion
    validator calling a per-state validity function once per interpolated state (reference
    src/ramp/motion_planner.py:366,469-483,612-625) -- bounds, then one collision check of one
    configuration per call.  Same interface as GpuMotionValidator, so the same planner drives both."""

    def __init__(self, planner, desc, qmap):
        self.p, self.desc, self.qmap = planner, desc, qmap
        self.valid_count = self.invalid_count = self.states_checked = self.states_offered = 0

    def state_valid(self, s) -> bool:
        import numpy as np

        from oracle import oracle
        from ramp_b200 import spaces

        self.states_checked += 1
        if not spaces.satisfies_bounds(self.p.space, s):
            return False
        q = self.qmap.apply(np.asarray(s, dtype=np.float64)[None])
        return not bool(oracle.check(self.desc, q, mode=oracle.MODE_STOP_AT_FIRST, want_pair_dist=False, nthreads=1)["hit"][0])

    def checkMotion(self, s1, s2, last_valid=None) -> bool:  # noqa: N802
        from ramp_b200 import spaces

        nd = spaces.valid_segment_count(self.p.space, s1, s2)
        ok, j = self.state_valid(s2), 1  # OMPL tests the end state first, then the interior ones in order
        while ok and j < nd:
            ok = self.state_valid(spaces.interpolate(self.p.space, s1, s2, j / nd))
            j += 1
        self.valid_count += ok
        self.invalid_count += not ok
        self.states_offered += max(nd, 1)
        return ok

    def check_edges(self, q0, q1) -> dict:
        import torch

        return {"valid": torch.tensor([self.checkMotion(a, b) for a, b in zip(q0, q1)])}


def measure_next_rows(dev, with_cpu: bool) -> dict:
    """The rows SURVEY 8(f) ranks next, measured on the Panda stand-in: f1 all-pairs sampling for the
    ALWAYS / NEVER reasons, f3 clearance (smallest signed distance), f4 batched differential-IK seeds
    (f2 is the FK and edge rows above).  Device-resident inputs; the CPU column is the oracle
    (OpenMP, all host threads) on a bounded sample."""
    import numpy as np
    import torch

    from ramp_b200 import RobotState, workloads
    from ramp_b200.desc import ModelDesc

    rm, _ = workloads.build("panda_64m")
    gpu = rm.gpu
    # the headline leg leaves ~1 GB of queue blocks in this model's pool; the pool hands what exceeds its
    # retention limit back to the driver at some later synchronisation (a 25 ms stall inside whichever
    # small call is being timed then): hand it all back now
    gpu.trim()
    desc = ModelDesc(rm.model, rm.collision_model)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def timed(fn, steps):
        fn()
        torch.cuda.synchronize()
        a.record()
        for _ in range(steps):
            fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) * 1e-3 / steps

    out = {"workload": "next_rows", "robot": "panda/configs", "scene": "capsule", "model": "stand-in"}
    n1 = 1 << 20
    q = workloads.sample_q_torch(rm.model, n1, seed=5, device=dev)
    dt = timed(lambda: gpu.check(q, pair_mask=True), 5)
    out["f1_all_pairs_mask"] = {"configs": n1, "pairs": len(rm.collision_model.collisionPairs), "configs_per_s": n1 / dt,
                                "api": "GpuModel.check(pair_mask=True): every pair of every configuration, one bit each"}
    n3 = 1 << 18
    dt = timed(lambda: gpu.clearance(q[:n3]), 2)
    out["f3_clearance"] = {"configs": n3, "configs_per_s": n3 / dt,
                           "api": "GpuModel.clearance (the d_min_dist output of ramp_check_batch): smallest signed distance over the pairs, GJK / EPA to convergence for the pairs whose capsule bounds can still lower it"}
    rs = RobotState.from_named_state(rm, "arm", "home")
    params = rm.ik_params("arm")
    n4 = 1 << 16
    rng = np.random.default_rng(4)
    lo, hi = np.maximum(params.lower, -3.0), np.minimum(params.upper, 3.0)
    goal = rng.uniform(lo, hi, (n4, len(lo)))
    seeds = np.clip(goal + rng.normal(scale=0.3, size=goal.shape), params.lower, params.upper).astype(np.float32)
    frame = rm.model.frames[rm.model.getFrameId(rm["arm"].tcp_link_name)]
    from ramp_b200 import spaces
    gpu.set_qmap(spaces.make_qmap(rm, "arm", rs.qpos))
    omi = gpu.forward_kinematics(torch.from_numpy(goal.astype(np.float32)).to(dev), actuated=True)[:, frame.parentJoint - 1]
    pl = torch.from_numpy(frame.placement.rt34().astype(np.float32)).to(dev)
    targets = torch.cat([omi[:, :, :3] @ pl[:, :3], (omi[:, :, :3] @ pl[:, 3] + omi[:, :, 3]).unsqueeze(2)], dim=2).contiguous()
    d_seeds = torch.from_numpy(seeds).to(dev)
    res = rm.differential_ik_batch("arm", d_seeds, targets, reference_state=rs)
    # (this call lasts as long as its slowest seed -- 500 dependent iterations of one warp -- and is
    # the one number of the run that moves between 5 and 17 ms with whatever else the device just did:
    # five calls timed one by one, median reported, fastest next to it)
    each = sorted(timed(lambda: rm.differential_ik_batch("arm", d_seeds, targets, reference_state=rs), 1) for _ in range(5))
    dt = each[2]
    out["f4_ik_seeds"] = {"seeds": n4, "seeds_per_s": n4 / dt, "converged_fraction": float(res["ok"].float().mean().item()),
                          "mean_iterations": float(res["iters"].float().mean().item()), "max_iterations": int(params.struct.max_iterations),
                          "seeds_per_s_fastest_call": n4 / each[0],
                          "api": "RobotModel.differential_ik_batch (ramp_ik_batch): damped least squares to 1e-3, one lane per seed"}
    # a goal-state generator keeps the seeds that converge and has no use for the last per cent that
    # would run to the reference's 500 iterations: the same call with 64
    res64 = rm.differential_ik_batch("arm", d_seeds, targets, reference_state=rs, max_iterations=64)
    dt64 = timed(lambda: rm.differential_ik_batch("arm", d_seeds, targets, reference_state=rs, max_iterations=64), 3)
    out["f4_ik_seeds"]["max_iterations_64"] = {"seeds_per_s": n4 / dt64, "converged_fraction": float(res64["ok"].float().mean().item())}
    if with_cpu:
        from oracle import oracle

        threads = os.cpu_count() or 1
        qs = q[: 1 << 15].cpu().numpy().astype(np.float64)
        t0 = time.perf_counter()
        oracle.check(desc, qs, mode=0, want_pair_dist=False, nthreads=threads)
        out["f1_all_pairs_mask"]["cpu_oracle"] = {"configs_per_s": len(qs) / (time.perf_counter() - t0), "cores": threads, "sample": len(qs)}
        qs = qs[: 1 << 13]
        t0 = time.perf_counter()
        oracle.check(desc, qs, nthreads=threads)
        out["f3_clearance"]["cpu_oracle"] = {"configs_per_s": len(qs) / (time.perf_counter() - t0), "cores": threads, "sample": len(qs)}
        k = 1 << 12
        qmap = spaces.make_qmap(rm, "arm", rs.qpos)
        t0 = time.perf_counter()
        oracle.ik(desc, qmap, params, seeds[:k].astype(np.float64), targets[:k].cpu().numpy().astype(np.float64))
        out["f4_ik_seeds"]["cpu_oracle"] = {"seeds_per_s": k / (time.perf_counter() - t0), "cores": threads, "sample": k}
    return out


def planning_leg(seeds: int, cpu_seeds: int) -> dict:
    """BASELINE config 3 as SURVEY 8(d) specifies it: FR3 (stand-in), start ``home``, goal
    [0, 0.25, 0, -1.25, 0, 1.5, 0] (reference examples/motion_planning.py:27), the reference capsule
    plus two plates that leave the goal pose 5 mm of clearance above and below (a narrow slot: the
    straight edge is invalid and RRTConnect needs ~100 edges), 5 s timeout, ``seeds`` seeds through
    the batched GPU validator (one ramp_check_edges_host call per edge) and ``cpu_seeds`` of the same
    seeds through the per-state CPU validator."""
    import numpy as np

    from ramp_b200 import MotionPlanner, RobotState, spaces, workloads
    from ramp_b200.desc import ModelDesc
    from ramp_b200.robot_model import create_geometry_object
    from ramp_b200.se3 import SE3
    from ramp_b200.shapes import Box

    rm, rs = workloads.build("fr3_edges")
    for k, z in enumerate((0.598, 0.803)):
        rs.add_object(create_geometry_object(f"plate{k}", Box(0.5, 0.8, 0.02), SE3(np.eye(3), np.asarray([0.65, 0.0, z]))))
    planner = MotionPlanner(rm, "arm")
    start = RobotState.from_named_state(rm, "arm", "home")
    goal = [0.0, 0.25, 0.0, -1.25, 0.0, 1.5, 0.0]
    gpu_validator = planner.validator

    def run(validator, n_seeds):
        planner.validator = validator
        rows = []
        for seed in range(n_seeds):
            planner.rng = np.random.default_rng(seed)
            validator.valid_count = validator.invalid_count = validator.states_checked = validator.states_offered = 0
            t0 = time.perf_counter()
            path = planner.plan(start, goal, timeout=5.0)
            dt = time.perf_counter() - t0
            # states of the edges offered (nd each): the GPU checks all of them per call, the per-state
            # validator stops at an edge's first invalid state
            rows.append((path is not None, dt, validator.valid_count + validator.invalid_count,
                         getattr(validator, "states_offered", 0) or validator.states_checked))
        if not rows:
            return None
        f, t, e, st = (np.asarray(c, dtype=np.float64) for c in zip(*rows))
        return {"seeds": n_seeds, "found": int(f.sum()), "plan_seconds_mean": float(t.mean()), "plan_seconds_max": float(t.max()),
                "edges_checked_mean": float(e.mean()), "states_checked_mean": float(st.mean()),
                "mean_nd": float(st.sum() / max(e.sum(), 1)), "edges_per_s": float(e.sum() / t.sum()),
                "states_per_s": float(st.sum() / t.sum())}

    planner.plan(start, goal, timeout=5.0)  # warm-up: library load, qmap upload
    gpu = run(gpu_validator, seeds)
    planner.set_reference_state(start)
    cpu = run(CpuStateValidator(planner, ModelDesc(rm.model, rm.collision_model), planner._qmap), cpu_seeds) if cpu_seeds > 0 else None
    planner.validator = gpu_validator
    out = {"planner": "built-in RRTConnect + shortcutting + path check (OMPL is not installable here)",
           "scene": "reference capsule + two plates forming a slot around the goal pose", "timeout_s": 5.0,
           "gpu_batched_validator": gpu, "cpu_per_state_validator": cpu,
           "note": "plan time includes the Python planner (nearest-neighbour search, tree bookkeeping), identical in both columns"}
    if gpu and cpu:
        out["plan_time_ratio_cpu_over_gpu"] = cpu["plan_seconds_mean"] / gpu["plan_seconds_mean"]
    return out


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="", help=f"headline workload (default {HEADLINE})")
    ap.add_argument("--n", "--rows", dest="n", type=int, default=0,
                    help="configurations per pass per GPU (--rows under torchrun, whose own parser claims --n)")
    ap.add_argument("--passes", type=int, default=0, help="calls per step (default: enough for a 25 ms step)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-subs", action="store_true", help="headline only (no `workloads` array)")
    ap.add_argument("--kernel", default="check", choices=["check", "fk"], help="fk: time ramp_fk_batch (poses stored) alone")
    ap.add_argument("--plan-seeds", type=int, default=20)
    args = ap.parse_args()

    # The CPU arm branches here, before anything of the GPU library is built, imported or loaded.
    if args.impl == "reference":
        run_reference(args)
        return
    args.warmup = max(args.warmup, 3)

    import __graft_entry__

    rank = int(os.environ.get("RANK", "0"))
    if rank == 0:
        __graft_entry__.build()

    import torch
    import torch.distributed as dist

    from ramp_b200 import workloads

    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU fallback); use --impl reference for the CPU path")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
        dist.barrier()  # rank 0 has built the library before anyone loads it

    name = args.workload or HEADLINE
    if name == "fr3_edges":
        line = measure_edges(name, args.n or workloads.WORKLOADS[name].n, args.steps, args.warmup, dev, seeds=args.plan_seeds, cpu_queries=args.plan_seeds)
        print(json.dumps({**line, "n_gpus": 1, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
                          "data": "synthetic"}))
        return
    if args.kernel == "fk":
        line = measure_fk(name, args.n or min(workloads.WORKLOADS[name].n, 1 << 22), args.steps, args.warmup, dev)
        print(json.dumps({**line, "n_gpus": 1, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
                          "data": "synthetic"}))
        return

    n = args.n or workloads.WORKLOADS[name].n
    passes = plan_passes(name, n, args.passes)
    head = measure_check(name, n, passes, args.steps, args.warmup, world=world, rank=rank, dev=dev,
                         cpu_seconds=0.0 if (args.no_cpu_baseline or world > 1) else 12.0, e2e_steps=max(3, min(args.steps, 5)))
    subs = []
    if world == 1 and not args.no_subs and not args.workload:
        torch.cuda.empty_cache()
        for sub, sn in SUBS:
            subs.append(measure_check(sub, sn, plan_passes(sub, sn, 0), 5, 3, world=1, rank=0, dev=dev,
                                      cpu_seconds=0.0 if args.no_cpu_baseline else 2.0, e2e_steps=3))
            torch.cuda.empty_cache()
        subs.append(measure_fk("panda_64m", 1 << 22, 10, 3, dev))
        # (before the planning leg, whose per-state CPU validator sets the oracle's OpenMP team to one thread)
        subs.append(measure_next_rows(dev, not args.no_cpu_baseline))
        subs.append(measure_edges("fr3_edges", 1 << 14, 5, 3, dev, seeds=args.plan_seeds, cpu_queries=0 if args.no_cpu_baseline else args.plan_seeds))
    if rank == 0:
        out = {"metric": METRIC, "value": head["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
               "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
               "dtype": "f32", "data": "synthetic"}
        out.update({k: v for k, v in head.items() if k not in ("workload", "metric", "unit", "value", "ms_per_step", "steps", "warmup")})
        if subs:
            out["workloads"] = subs
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
