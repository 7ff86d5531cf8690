"""ramp_check_host on batches of one slice: how the split into sections (one queue kernel each) moves the
end-to-end time.  RAMP_B200_SECTIONS is read per call.  python profiles/e2e_sections.py"""
import json, os, sys, time
sys.path.insert(0, ".")
import numpy as np, torch
from ramp_b200 import workloads
from ramp_b200.engine import pinned_empty
settings = ["", "0.34,0.33,0.33", "0.4,0.35,0.25", "0.45,0.35,0.2", "0.5,0.3,0.2", "0.4,0.3,0.2,0.1", "0.35,0.3,0.2,0.15",
            "0.5,0.25,0.15,0.1", "0.3,0.25,0.2,0.15,0.1", "0.6,0.4", "0.7,0.3", "0.55,0.3,0.15", "0.5,0.3,0.14,0.06"]
for name, n in (("ur5e_1m", 1 << 20), ("panda_64m", 1 << 20), ("ur5e_1m", 1 << 21), ("h1_16m", 1 << 20)):
    rm, _ = workloads.build(name)
    gpu = rm.gpu
    nq = rm.model.nq
    qd = workloads.sample_q_torch(rm.model, n, seed=1, device=gpu.device)
    qh = pinned_empty((n, nq), np.float32)
    torch.from_numpy(qh).copy_(qd[:, :nq])
    hh = pinned_empty((n,), np.uint8)
    ref = gpu.check(qd)["hit"].cpu().numpy()
    for sset in settings:
        if sset: os.environ["RAMP_B200_SECTIONS"] = sset
        else: os.environ.pop("RAMP_B200_SECTIONS", None)
        for _ in range(5): gpu.check_host_into(qh, hh)
        best = []
        for rep in range(3):
            t0 = time.perf_counter()
            for _ in range(20): gpu.check_host_into(qh, hh)
            best.append((time.perf_counter() - t0) / 20)
        dt = sorted(best)[1]
        print(json.dumps({"workload": name, "n": n, "sections": sset or "default", "ms": round(dt * 1e3, 4), "configs_per_s": n / dt,
                          "same": bool(np.array_equal(hh.astype(bool), ref))}), flush=True)
