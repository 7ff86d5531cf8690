"""Latency of one validity call against the batch size (device-resident rows, CUDA events over 200 calls),
and of one ramp_check_host call (host buffers).  Environment: RAMP_B200_SMALL_ROWS=<n> moves the switch
between the warp-per-configuration kernel and the tile kernel (experiments).
    python profiles/latency_probe.py [workload ...]"""
import json, os, sys, time
sys.path.insert(0, ".")
import numpy as np, torch
from ramp_b200 import workloads
from ramp_b200.engine import pinned_empty

names = sys.argv[1:] or ["panda_4096", "ur5e_1m", "h1_16m"]
for name in names:
    rm, _ = workloads.build(name)
    gpu = rm.gpu
    nq = rm.model.nq
    for n in (32, 256, 1024, 2048, 4096, 8192, 16384, 32768, 65536, 131072, 262144):
        qd = workloads.sample_q_torch(rm.model, n, seed=1, device=gpu.device)
        hit = torch.empty(n, dtype=torch.uint8, device=gpu.device)
        for _ in range(20): gpu.check_raw(qd, hit)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(200): gpu.check_raw(qd, hit)
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / 200
        # one blocking call at a time (what a planner sees): wall clock around call + synchronize
        t0 = time.perf_counter()
        for _ in range(100):
            gpu.check_raw(qd, hit)
            torch.cuda.synchronize()
        us_sync = (time.perf_counter() - t0) * 1e6 / 100
        qh = pinned_empty((n, nq), np.float32)
        torch.from_numpy(qh).copy_(qd[:, :nq])
        hh = pinned_empty((n,), np.uint8)
        for _ in range(5): gpu.check_host_into(qh, hh)
        t0 = time.perf_counter()
        for _ in range(100): gpu.check_host_into(qh, hh)
        us_host = (time.perf_counter() - t0) * 1e6 / 100
        print(json.dumps({"workload": name, "n": n, "small_rows": os.environ.get("RAMP_B200_SMALL_ROWS", "default"),
                          "us_per_call_stream": round(us, 2), "us_per_call_blocking": round(us_sync, 2),
                          "us_per_host_call": round(us_host, 2), "hit_sum": int(hit.sum().item())}), flush=True)
