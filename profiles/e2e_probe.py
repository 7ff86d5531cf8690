"""End-to-end probe of ramp_check_host: configurations/s and H2D GB/s per workload and batch size.
    python profiles/e2e_probe.py [wc]      (wc: input buffer in write-combined page-locked memory)"""
import json, os, sys, time
sys.path.insert(0, ".")
wc = len(sys.argv) > 1 and sys.argv[1] == "wc"
import numpy as np, torch
from ramp_b200 import workloads
from ramp_b200.engine import pinned_empty
for name, n in (("panda_64m", 1 << 24), ("ur5e_1m", 1 << 20), ("ur5e_1m", 1 << 24), ("h1_16m", 1 << 22)):
    rm, _ = workloads.build(name)
    gpu = rm.gpu
    nq = rm.model.nq
    qd = workloads.sample_q_torch(rm.model, n, seed=1, device=gpu.device)
    os.environ["RAMP_B200_HOST_WC"] = "1" if wc else "0"
    qh = pinned_empty((n, nq), np.float32)
    os.environ["RAMP_B200_HOST_WC"] = "0"
    torch.from_numpy(qh).copy_(qd[:, :nq])
    hh = pinned_empty((n,), np.uint8)
    gpu.check_host_into(qh, hh)
    same = bool(np.array_equal(hh.astype(bool), gpu.check(qd)["hit"].cpu().numpy()))
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(10): gpu.check_host_into(qh, hh)
    dt = (time.perf_counter() - t0) / 10
    print(json.dumps({"workload": name, "n": n, "write_combined_input": wc, "e2e_configs_per_s": n / dt, "ms": dt * 1e3,
                      "h2d_gbs": n * nq * 4 / dt / 1e9, "same": same}))
