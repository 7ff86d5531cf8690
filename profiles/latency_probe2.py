"""Where a small validity call spends its time: CPU time to enqueue one call against the GPU time
between its start and end (CUDA events), per batch size.  python profiles/latency_probe2.py [iters]"""
import json, os, sys, time
sys.path.insert(0, ".")
import torch
from ramp_b200 import workloads
iters = int(sys.argv[1]) if len(sys.argv) > 1 else 200
for name in ("panda_4096", "h1_16m"):
    rm, _ = workloads.build(name)
    gpu = rm.gpu
    for n in (32, 1024, 4096, 16384):
        qd = workloads.sample_q_torch(rm.model, n, seed=1, device=gpu.device)
        hit = torch.empty(n, dtype=torch.uint8, device=gpu.device)
        for _ in range(10): gpu.check_raw(qd, hit)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        for _ in range(iters): gpu.check_raw(qd, hit)
        e1.record()
        t1 = time.perf_counter()
        torch.cuda.synchronize()
        # a single call, GPU idle before it: the kernel's own duration plus launch latency
        single = []
        for _ in range(10):
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); gpu.check_raw(qd, hit); b.record(); torch.cuda.synchronize()
            single.append(a.elapsed_time(b) * 1e3)
        print(json.dumps({"workload": name, "n": n, "small_rows": os.environ.get("RAMP_B200_SMALL_ROWS", "default"),
                          "cpu_enqueue_us": round((t1 - t0) * 1e6 / iters, 2), "gpu_us_per_call_back_to_back": round(e0.elapsed_time(e1) * 1e3 / iters, 2),
                          "gpu_us_single_call_median": round(sorted(single)[5], 2)}), flush=True)
