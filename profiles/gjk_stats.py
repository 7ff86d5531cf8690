"""GJK work-queue statistics per workload (items per configuration, iterations per item, lane
occupancy of the simplex trips), from a library built with -DRAMP_STATS:

    nvcc ... -DRAMP_STATS -o ramp_b200/libramp_b200_stats.so ramp_b200/csrc/ramp_b200.cu   (here)
    RAMP_B200_LIB=ramp_b200/libramp_b200_stats.so python profiles/gjk_stats.py [workload ...]   (GPU box)
"""
import ctypes as C
import json
import sys

import torch

sys.path.insert(0, ".")
from ramp_b200 import engine, workloads  # noqa: E402

names = sys.argv[1:] or ["ur5e_1m", "panda_64m", "panda_boxy_64m", "h1_16m", "h1_boxy_16m"]
lib = engine.load_library()
n = 1 << 20
for name in names:
    rm, _ = workloads.build(name)
    gpu = rm.gpu
    q = workloads.sample_q_torch(rm.model, n, seed=0xB200, device=gpu.device)
    hit = torch.empty(n, dtype=torch.uint8, device=gpu.device)
    gpu.check_raw(q, hit)
    st = (C.c_ulonglong * 64)()
    lib.ramp_debug_stats(st, 1)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ev[0].record()
    for _ in range(5):
        gpu.check_raw(q, hit)
    ev[1].record()
    torch.cuda.synchronize()
    lib.ramp_debug_stats(st, 1)
    s = [int(x) / 5 for x in st]
    items, iters, hits, skipped, started, lanes = s[0], s[1], s[2], s[3], s[4], s[5]
    hist = {i: s[8 + i] for i in range(41) if s[8 + i] > 0}
    print(json.dumps({"workload": name, "n": n, "colliding": float(hit.float().mean()),
                      "items_per_config": (started + skipped) / n, "started_per_config": started / n, "skipped_per_config": skipped / n,
                      "iterations_per_item": iters / max(items, 1), "hit_fraction_of_items": hits / max(items, 1),
                      "active_lanes_per_simplex_trip": lanes / max(iters, 1),
                      "simplex_size_share": {k: s[56 + k] / max(iters, 1) for k in range(5)},
                      "iterations_histogram": {k: round(v / max(items, 1), 4) for k, v in hist.items()},
                      "ms_per_call_with_stats_atomics": ev[0].elapsed_time(ev[1]) / 5}))
