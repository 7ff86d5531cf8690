"""A parity sweep much larger than the test suite's: per robot, N random configurations through
ramp_check_batch (flags, first pair, every pair's bit) against the FP64 oracle -- configurations and
pairs outside the +-1e-4 m band must agree bit for bit.  One JSON line per robot.
    python profiles/parity_sweep.py [N]         (default 500 000; H1 robots a fifth of that)"""
import json, sys, time
sys.path.insert(0, ".")
sys.path.insert(0, "tests")
import numpy as np, torch
import __graft_entry__
__graft_entry__.build()
from oracle import oracle
from ramp_b200 import workloads, load_robot_model, RobotState
from ramp_b200.desc import ModelDesc
import test_gpu_parity as T

BAND = 1e-4
N = int(sys.argv[1]) if len(sys.argv) > 1 else 500_000
cases = [("panda/configs", "capsule", N), ("ur5e/configs", "obstacles", N), ("fr3/configs", "capsule", N),
         ("panda_boxy/configs", "capsule", N), ("unitree_h1/configs", None, N // 5), ("unitree_h1_boxy/configs", None, N // 5),
         ("joint_zoo", "capsule", N), ("rrr/floating_configs", "capsule", N), ("floating", "obstacles", N)]
for name, scene, n in cases:
    rm, _ = T._setup(name, scene)
    d = ModelDesc(rm.model, rm.collision_model)
    tot = dict(robot=name, scene=scene, n=0, colliding=0, in_band=0, bad_cfg=0, bad_pairs=0, bad_first=0, bad_early_out=0)
    t0 = time.time()
    for lo in range(0, n, 100_000):
        m = min(100_000, n - lo)
        q = workloads.sample_q(rm.model, m, seed=1000 + lo).astype(np.float32)
        ref = oracle.check(d, q[:, : rm.model.nq].astype(np.float64))
        qd = torch.from_numpy(q).cuda()
        res = rm.gpu.check(qd, first_pair=True, pair_mask=True)
        hit, first, mask = res["hit"].cpu().numpy(), res["first_pair"].cpu().numpy(), res["pair_mask"].cpu().numpy()
        early = rm.gpu.check(qd)["hit"].cpu().numpy()
        pd = ref["pair_dist"]
        outside = np.abs(ref["min_dist"]) > BAND
        pair_out = np.abs(pd) > BAND
        clean = pair_out.all(axis=1)
        tot["n"] += m; tot["colliding"] += int(ref["hit"].sum()); tot["in_band"] += int((~outside).sum())
        tot["bad_cfg"] += int((hit != ref["hit"])[outside].sum())
        tot["bad_pairs"] += int(((mask != (pd <= 1e-12)) & pair_out).sum())
        tot["bad_first"] += int((first != ref["first_pair"])[clean].sum())
        tot["bad_early_out"] += int((early != ref["hit"])[outside].sum())
    tot["seconds"] = round(time.time() - t0, 1)
    print(json.dumps(tot), flush=True)
