import sys, time, json
sys.path.insert(0, ".")
import numpy as np, torch
from ramp_b200 import workloads
from ramp_b200.engine import pinned_empty
for name, n in (("ur5e_1m", 1 << 20), ("panda_64m", 1 << 20), ("panda_64m", 1 << 24), ("h1_16m", 1 << 22), ("ur5e_1m", 1 << 24), ("ur5e_1m", 300001)):
    rm, _ = workloads.build(name)
    gpu = rm.gpu
    nq = rm.model.nq
    q = workloads.sample_q(rm.model, n, seed=1)
    qh = pinned_empty((n, nq), np.float32); qh[:] = q[:, :nq]
    hh = pinned_empty((n,), np.uint8)
    gpu.check_host_into(qh, hh)
    ref = gpu.check(torch.from_numpy(q).cuda())["hit"].cpu().numpy()
    same = bool(np.array_equal(hh.astype(bool), ref))
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(10): gpu.check_host_into(qh, hh)
    dt = (time.perf_counter() - t0) / 10
    print(json.dumps({"workload": name, "n": n, "e2e_configs_per_s": n / dt, "ms": dt * 1e3, "h2d_gbs": n * nq * 4 / dt / 1e9, "same": same}))
