"""A parity sweep much larger than the test suite's: per robot, N random configurations through
ramp_check_batch (flags, first pair, every pair's bit) against the FP64 oracle -- configurations and
pairs outside the +-1e-4 m band must agree bit for bit.  One JSON line per robot.
    python profiles/parity_sweep.py [N]         (default 500 000; H1 robots a fifth of that)
    python profiles/parity_sweep.py N more      also: signed distances and joint poses on N / 5 configurations, planner
                                                edges (N / 25) through ramp_check_edges, the small robots of the zoo"""
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
MORE = len(sys.argv) > 2 and sys.argv[2] == "more"
cases = [("panda/configs", "capsule", N), ("ur5e/configs", "obstacles", N), ("fr3/configs", "capsule", N),
         ("panda_boxy/configs", "capsule", N), ("unitree_h1/configs", None, N // 5), ("unitree_h1_boxy/configs", None, N // 5),
         ("joint_zoo", "capsule", N), ("rrr/floating_configs", "capsule", N), ("floating", "obstacles", N)]
if MORE:
    cases += [("rrr/configs", None, N), ("rrr/planar_configs", "capsule", N), ("rrr_mj/configs", "obstacles", N), ("planar_rrr/configs", "capsule", N)]
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


if MORE:
    from ramp_b200 import spaces
    for name, scene, n in cases:
        rm, _ = T._setup(name, scene)
        d = ModelDesc(rm.model, rm.collision_model)
        # ---- signed distances of every pair and joint poses
        m = max(n // 5, 1000)
        tot = dict(robot=name, what="distances+poses", n=0, max_pose_err=0.0, max_sep_err=0.0, max_pen_err=0.0, sign_flips_outside_band=0)
        for lo in range(0, m, 50_000):
            k = min(50_000, m - lo)
            q = workloads.sample_q(rm.model, k, seed=7000 + lo).astype(np.float32)
            q64 = q[:, : rm.model.nq].astype(np.float64)
            qd = torch.from_numpy(q).cuda()
            ref = oracle.check(d, q64)["pair_dist"]
            got = rm.gpu.distance(qd).cpu().numpy().astype(np.float64)
            sep, pen = ref > BAND, ref < -BAND
            if sep.any(): tot["max_sep_err"] = max(tot["max_sep_err"], float((np.abs(got - ref) / (1.0 + np.abs(ref)))[sep].max()))
            if pen.any(): tot["max_pen_err"] = max(tot["max_pen_err"], float((np.abs(got - ref) / (1e-3 + 0.025 * np.abs(ref)))[pen].max()))
            tot["sign_flips_outside_band"] += int((((got > 0) != (ref > 0)) & (np.abs(ref) > BAND)).sum())
            fk = rm.gpu.forward_kinematics(qd).cpu().numpy()
            tot["max_pose_err"] = max(tot["max_pose_err"], float(np.abs(fk - oracle.fk(d, q64)).max()))
            tot["n"] += k
        print(json.dumps(tot), flush=True)
        # ---- the one-warp-per-configuration kernel (calls of <= 8000 rows), same comparison
        tot = dict(robot=name, what="small-batch path", n=0, in_band=0, bad_cfg=0, bad_pairs=0, bad_first=0)
        ms = max(n // 5, 8000)
        for lo in range(0, ms, 80_000):
            k = min(80_000, ms - lo)
            q = workloads.sample_q(rm.model, k, seed=9000 + lo).astype(np.float32)
            ref = oracle.check(d, q[:, : rm.model.nq].astype(np.float64))
            hit = np.empty(k, dtype=bool); first = np.empty(k, dtype=np.int32); mask = np.empty((k, len(rm.collision_model.collisionPairs)), dtype=bool)
            for c0 in range(0, k, 2000):
                res = rm.gpu.check(torch.from_numpy(q[c0:c0 + 2000]).cuda(), first_pair=True, pair_mask=True)
                hit[c0:c0 + 2000], first[c0:c0 + 2000], mask[c0:c0 + 2000] = res["hit"].cpu().numpy(), res["first_pair"].cpu().numpy(), res["pair_mask"].cpu().numpy()
            pd = ref["pair_dist"]
            outside = np.abs(ref["min_dist"]) > BAND
            pair_out = np.abs(pd) > BAND
            tot["n"] += k; tot["in_band"] += int((~outside).sum())
            tot["bad_cfg"] += int((hit != ref["hit"])[outside].sum())
            tot["bad_pairs"] += int(((mask != (pd <= 1e-12)) & pair_out).sum())
            tot["bad_first"] += int((first != ref["first_pair"])[pair_out.all(axis=1)].sum())
        print(json.dumps(tot), flush=True)
        # ---- planner edges
        if "arm" not in rm.groups:
            continue
        space = spaces.make_space(rm, "arm"); qmap = spaces.make_qmap(rm, "arm")
        rm.gpu.set_qmap(qmap)
        ne = max(n // 25, 1000)
        rng = np.random.default_rng(3)
        q0 = spaces.sample_uniform(space, ne, rng).astype(np.float32)
        q1 = (q0 + rng.normal(scale=0.3, size=q0.shape)).astype(np.float32)
        q1s = spaces.sample_uniform(space, ne, rng).astype(np.float32)
        for sp in space.subspaces:
            c = sp["col"]
            if sp["type"] in (3, 4):
                o = 3 if sp["type"] == 3 else 0
                q1[:, c + o : c + o + 4] = q1s[:, c + o : c + o + 4]
            if sp["type"] in (1, 2):
                kk = c + (2 if sp["type"] == 2 else 0)
                q1[:, kk] = q1s[:, kk]
        # ---- the same verdicts from ACTUATED values of the group (q map: reference state, mimic joints, continuous joints)
        na = max(n // 5, 1000)
        act = spaces.sample_uniform(space, na, np.random.default_rng(11)).astype(np.float32)
        refa = oracle.check(d, oracle.qmap_apply(qmap, act.astype(np.float64)))
        resa = rm.gpu.check(torch.from_numpy(act).cuda(), actuated=True, first_pair=True, pair_mask=True)
        pda = refa["pair_dist"]; outa = np.abs(refa["min_dist"]) > BAND
        print(json.dumps(dict(robot=name, what="actuated rows", n=na, in_band=int((~outa).sum()),
                              bad_cfg=int((resa["hit"].cpu().numpy() != refa["hit"])[outa].sum()),
                              bad_pairs=int(((resa["pair_mask"].cpu().numpy() != (pda <= 1e-12)) & (np.abs(pda) > BAND)).sum()),
                              bad_first=int((resa["first_pair"].cpu().numpy() != refa["first_pair"])[(np.abs(pda) > BAND).all(axis=1)].sum()))), flush=True)
        got = {k_: v.cpu().numpy() for k_, v in rm.gpu.check_edges(torch.from_numpy(q0).cuda(), torch.from_numpy(q1).cuda(), space).items()}
        ref = oracle.check_edges(d, qmap, space, q0.astype(np.float64), q1.astype(np.float64))
        same_nd = got["nd"] == ref["nd"]
        agree = (got["valid"] == ref["valid"]) & (got["first_invalid"] == ref["first_invalid"])
        bad = np.flatnonzero(~agree & same_nd)
        unexplained = 0
        for e in bad:
            n_d = int(ref["nd"][e])
            js = {int(got["first_invalid"][e]), int(ref["first_invalid"][e])} - {-1}
            ok = False
            for j in js:
                st = q1[e].astype(np.float64) if j >= n_d else oracle.edge_interpolate(space, q0[e], q1[e], j / n_d)
                ok = ok or abs(oracle.check(d, oracle.qmap_apply(qmap, st))["min_dist"][0]) <= BAND
            unexplained += not ok
        print(json.dumps(dict(robot=name, what="edges", edges=ne, invalid=int((~ref["valid"]).sum()), nd_off_by_one=int((~same_nd).sum()),
                              nd_off_by_more=int((np.abs(got["nd"] - ref["nd"]) > 1).sum()), disagreements_in_band=len(bad) - unexplained,
                              unexplained=unexplained)), flush=True)
