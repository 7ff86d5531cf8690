"""GPU (FP32, CUDA) vs CPU oracle (FP64) on the same seeded inputs, through the C ABI.

Tolerances (BASELINE.json north_star): link poses within 1e-5 (m for translations; rotation
entries, i.e. < 1e-5 rad), collision booleans bit-exact for every configuration whose oracle
signed distance lies outside +-1e-4 m; the in-band count is reported.
"""

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

POSE_TOL = 1e-5
BAND = 1e-4

ZOO = ["rrr/configs", "rrr/planar_configs", "rrr/floating_configs", "planar_rrr/configs", "rrr_mj/configs",
       "acrobot/configs", "double_pendulum", "floating", "planar", "panda/configs", "fr3/configs", "ur5e/configs",
       "unitree_h1/configs", "joint_zoo"]


def _setup(name, scene=None):
    from ramp_b200 import workloads
    from ramp_b200.robot_model import load_robot_model
    from ramp_b200.robot_state import RobotState

    rm = load_robot_model("zoo::" + name)
    rs = RobotState(rm)
    if scene == "capsule":
        rs.add_object(workloads.reference_capsule())
    elif scene == "obstacles":
        for o in workloads.obstacle_scene():
            rs.add_object(o)
    return rm, rs


@pytest.mark.parametrize("name", ZOO)
def test_fk_matches_oracle(built, name):
    from oracle import oracle
    from ramp_b200 import workloads
    from ramp_b200.desc import ModelDesc

    rm, _ = _setup(name)
    n = 1000  # not a multiple of 32: exercises the ragged last tile
    q = workloads.sample_q(rm.model, n, seed=1)
    got = rm.gpu.forward_kinematics(torch.from_numpy(q).cuda()).cpu().numpy()
    ref = oracle.fk(ModelDesc(rm.model, rm.collision_model), q[:, : rm.model.nq].astype(np.float64))
    assert got.shape == ref.shape
    assert np.abs(got - ref).max() < POSE_TOL


@pytest.mark.parametrize("name", ["rrr/floating_configs", "panda/configs", "unitree_h1/configs"])
def test_placements_match_oracle(built, name):
    from oracle import oracle
    from ramp_b200 import workloads
    from ramp_b200.desc import ModelDesc

    rm, _ = _setup(name, "capsule")
    q = workloads.sample_q(rm.model, 333, seed=2)
    got = rm.gpu.placements(torch.from_numpy(q).cuda()).cpu().numpy()
    ref = oracle.placements(ModelDesc(rm.model, rm.collision_model), q[:, : rm.model.nq].astype(np.float64))
    assert np.abs(got - ref).max() < POSE_TOL


def _compare_check(rm, q, report):
    from oracle import oracle
    from ramp_b200.desc import ModelDesc

    d = ModelDesc(rm.model, rm.collision_model)
    ref = oracle.check(d, q[:, : rm.model.nq].astype(np.float64))
    res = rm.gpu.check(torch.from_numpy(q).cuda(), first_pair=True, pair_mask=True)
    hit = res["hit"].cpu().numpy()
    first = res["first_pair"].cpu().numpy()
    mask = res["pair_mask"].cpu().numpy()
    pd = ref["pair_dist"]
    outside = np.abs(ref["min_dist"]) > BAND
    # every pair decided on its own: bit-exact outside the band
    pair_out = np.abs(pd) > BAND
    ref_mask = pd <= 1e-12
    assert mask.shape == ref_mask.shape
    bad_pairs = int(((mask != ref_mask) & pair_out).sum())
    bad_cfg = int((hit != ref["hit"])[outside].sum())
    # first colliding pair (list order) wherever no pair of the configuration is in the band
    clean = pair_out.all(axis=1)
    bad_first = int((first != ref["first_pair"])[clean].sum())
    report(n=len(q), colliding=int(ref["hit"].sum()), in_band=int((~outside).sum()), bad_cfg=bad_cfg,
           bad_pairs=bad_pairs, bad_first=bad_first)
    assert bad_cfg == 0 and bad_pairs == 0 and bad_first == 0
    # early-out mode must give the same flags
    res2 = rm.gpu.check(torch.from_numpy(q).cuda())
    assert (res2["hit"].cpu().numpy() == hit)[outside].all()


CHECK_CASES = [("rrr/configs", None, 20000), ("rrr/planar_configs", "capsule", 20000),
               ("rrr/floating_configs", "capsule", 20000), ("rrr_mj/configs", "obstacles", 20000),
               ("planar_rrr/configs", "capsule", 20000), ("floating", "obstacles", 20000),
               ("panda/configs", "capsule", 20000), ("fr3/configs", "capsule", 20000),
               ("ur5e/configs", "obstacles", 20000), ("unitree_h1/configs", None, 4000),
               ("unitree_h1/configs", None, 1500),  # (177 pairs: one warp per configuration up to 2962 rows)
               # one joint of every type (RUB-unaligned, Spherical, SphericalZYX, Translation ... included),
               # on the tile path and on the one-warp-per-configuration path (n <= 8192 for an arm)
               ("joint_zoo", "capsule", 20000), ("joint_zoo", "capsule", 3000),
               # non-capsule link hulls: rounded boxes, elbows, plates, 100-vertex blobs
               ("panda_boxy/configs", "capsule", 20000), ("panda_boxy/configs", "capsule", 3000),
               ("unitree_h1_boxy/configs", None, 4000)]


@pytest.mark.parametrize("name,scene,n", CHECK_CASES)
def test_check_matches_oracle(built, name, scene, n, record_property):
    from ramp_b200 import workloads

    rm, _ = _setup(name, scene)
    q = workloads.sample_q(rm.model, n, seed=3)

    def report(**kw):
        print(name, kw)
        for k, v in kw.items():
            record_property(k, v)

    _compare_check(rm, q, report)


def test_kat_states_valid(built):
    """The reference's own known answers: home / goal states are collision free."""
    from ramp_b200.robot_model import as_pinocchio_qpos
    from ramp_b200.kinematics import neutral

    for cfg, scene, goal in (("rrr/configs", None, [0.0, 1.0, 1.0]), ("rrr/configs", "capsule", [0.0, 2.0, 1.4]),
                             ("rrr/planar_configs", "capsule", [1.0, -0.5, 1.57, 0.5, 0.25, 0.1]),
                             ("rrr/floating_configs", "capsule", [1.0, 0.5, 1.0, 1.0, 0.0, 0.0, 0.0, 0.5, 0.25, 0.1])):
        rm, rs = _setup(cfg, scene)
        home = as_pinocchio_qpos(rm, neutral(rm.model), "arm", rm["arm"].named_states["home"])
        g = as_pinocchio_qpos(rm, neutral(rm.model), "arm", goal)
        res = rm.gpu.check_host(np.stack([home, g]).astype(np.float32))
        assert not res["hit"].any(), (cfg, res)


def test_actuated_input_matches_host_map(built):
    """RAMP_Q_ACTUATED: mimic then (cos, sin) on the device == the host statement."""
    from oracle import oracle
    from ramp_b200 import spaces
    from ramp_b200.desc import ModelDesc

    for cfg, group in (("panda/configs", "arm"), ("double_pendulum", "arm"), ("rrr/planar_configs", "arm"),
                       ("planar_rrr/configs", "arm")):
        rm, _ = _setup(cfg)
        qmap = spaces.make_qmap(rm, group)
        rm.gpu.set_qmap(qmap)
        rng = np.random.default_rng(5)
        act = rng.uniform(-1.5, 1.5, (257, qmap.n_act)).astype(np.float32)
        got = rm.gpu.forward_kinematics(torch.from_numpy(act).cuda(), actuated=True).cpu().numpy()
        q = oracle.qmap_apply(qmap, act.astype(np.float64))
        assert np.allclose(q, qmap.apply(act.astype(np.float64)))
        ref = oracle.fk(ModelDesc(rm.model, rm.collision_model), q)
        assert np.abs(got - ref).max() < POSE_TOL


EDGE_CASES = [("rrr/configs", None, 20), ("panda/configs", "capsule", 32), ("rrr/floating_configs", "capsule", 7),
              ("rrr/configs", None, 300), ("rrr/planar_configs", "capsule", 200), ("rrr/floating_configs", "capsule", 150),
              ("panda/configs", "capsule", 200), ("planar_rrr/configs", "capsule", 200)]


@pytest.mark.parametrize("name,scene,n_edges", EDGE_CASES)
def test_edges_match_oracle(built, name, scene, n_edges):
    from oracle import oracle
    from ramp_b200 import spaces
    from ramp_b200.desc import ModelDesc

    rm, _ = _setup(name, scene)
    space = spaces.make_space(rm, "arm")
    qmap = spaces.make_qmap(rm, "arm")
    gpu = rm.gpu
    gpu.set_qmap(qmap)
    rng = np.random.default_rng(7)
    q0 = spaces.sample_uniform(space, n_edges, rng).astype(np.float32)
    q1 = (q0 + rng.normal(scale=0.35, size=q0.shape)).astype(np.float32)
    # keep quaternion parts unit and so2 parts in range
    q1s = spaces.sample_uniform(space, n_edges, rng).astype(np.float32)
    for s in space.subspaces:
        c = s["col"]
        if s["type"] in (3, 4):
            o = 3 if s["type"] == 3 else 0
            q1[:, c + o : c + o + 4] = q1s[:, c + o : c + o + 4]
        if s["type"] in (1, 2):
            k = c + (2 if s["type"] == 2 else 0)
            q1[:, k] = q1s[:, k]
    q1[0] = q0[0]  # a zero-length edge: nd = 0, only the end state is checked
    got = gpu.check_edges(torch.from_numpy(q0).cuda(), torch.from_numpy(q1).cuda(), space)
    d = ModelDesc(rm.model, rm.collision_model)
    ref = oracle.check_edges(d, qmap, space, q0.astype(np.float64), q1.astype(np.float64))
    nd = got["nd"].cpu().numpy()
    # FP32 vs FP64 ceil() may differ when dist/lvs sits on an integer: allow +-1 on few edges
    same_nd = nd == ref["nd"]
    assert (np.abs(nd - ref["nd"]) <= 1).all() and same_nd.mean() > (0.97 if n_edges > 100 else 0.8)
    valid = got["valid"].cpu().numpy()
    first = got["first_invalid"].cpu().numpy()
    agree = (valid == ref["valid"]) & (first == ref["first_invalid"])
    # disagreements must be explained by a near-contact state (margin band) or an nd tie
    bad = np.flatnonzero(~agree & same_nd)
    unexplained = 0
    for e in bad:
        n_d = int(ref["nd"][e])
        js = {int(first[e]), int(ref["first_invalid"][e])} - {-1}
        ok = False
        for j in js:
            st = q1[e].astype(np.float64) if j >= n_d else oracle.edge_interpolate(space, q0[e], q1[e], j / n_d)
            r = oracle.check(d, oracle.qmap_apply(qmap, st))
            ok = ok or abs(r["min_dist"][0]) <= BAND
        unexplained += not ok
    print(name, dict(edges=n_edges, invalid=int((~ref["valid"]).sum()), nd_mismatch=int((~same_nd).sum()),
                     disagreements=len(bad), unexplained=unexplained))
    assert unexplained == 0


def test_edges_large_batch_matches_small_batches(built):
    """A batch of more than one slice of edges (2^18; offsets scanned over 1000+ blocks of 256 edges)
    gives what the same edges give in batches of 3000: same segment counts, same verdicts and first
    invalid states but for margin-band states (the GJK of a short and of a long work queue share
    vertex scans differently), and a sample agrees with the oracle."""
    from oracle import oracle
    from ramp_b200 import spaces
    from ramp_b200.desc import ModelDesc

    rm, _ = _setup("panda/configs", "capsule")
    space = spaces.make_space(rm, "arm")
    qmap = spaces.make_qmap(rm, "arm")
    gpu = rm.gpu
    gpu.set_qmap(qmap)
    n_edges = (1 << 18) + 777
    rng = np.random.default_rng(11)
    q0 = spaces.sample_uniform(space, n_edges, rng).astype(np.float32)
    q1 = (q0 + rng.normal(scale=0.3, size=q0.shape)).astype(np.float32)
    d0, d1 = torch.from_numpy(q0).cuda(), torch.from_numpy(q1).cuda()
    big = {k: v.cpu().numpy() for k, v in gpu.check_edges(d0, d1, space).items()}
    valid, first, nd = (np.empty_like(big[k]) for k in ("valid", "first_invalid", "nd"))
    for lo in range(0, n_edges, 3000):
        part = gpu.check_edges(d0[lo:lo + 3000], d1[lo:lo + 3000], space)
        valid[lo:lo + 3000], first[lo:lo + 3000], nd[lo:lo + 3000] = (part[k].cpu().numpy() for k in ("valid", "first_invalid", "nd"))
    assert np.array_equal(big["nd"], nd)
    differ = np.flatnonzero((big["valid"] != valid) | (big["first_invalid"] != first))
    assert len(differ) <= 1e-4 * n_edges, f"{len(differ)} edges differ between one large and many small batches"
    assert 0.05 < big["valid"].mean() < 0.95
    d = ModelDesc(rm.model, rm.collision_model)
    for e in differ[:20]:  # each explained by a state in the margin band
        js = {int(big["first_invalid"][e]), int(first[e])} - {-1}
        n_d = int(nd[e])
        near = False
        for j in js:
            st = q1[e].astype(np.float64) if j >= n_d else oracle.edge_interpolate(space, q0[e], q1[e], j / n_d)
            near = near or abs(oracle.check(d, oracle.qmap_apply(qmap, st))["min_dist"][0]) <= BAND
        assert near
    pick = np.r_[np.arange(100), (1 << 18) - 50 + np.arange(100), n_edges - 100 + np.arange(100)]  # across the slice boundary
    ref = oracle.check_edges(d, qmap, space, q0[pick].astype(np.float64), q1[pick].astype(np.float64))
    same_nd = big["nd"][pick] == ref["nd"]
    agree = (big["valid"][pick] == ref["valid"]) & (big["first_invalid"][pick] == ref["first_invalid"])
    assert same_nd.mean() > 0.97 and (agree | ~same_nd).mean() > 0.98


def test_cases_found_by_the_parity_sweep(built):
    """The (configuration, pair) cases on which GPU and oracle disagreed outside the band in the first
    5M-per-robot sweep (tests/golden/sweep_cases.jsonl): shallow overlaps the FP32 GJK
    stalled on, a capsule / capsule pair with the non-rigid reference capsule, deep cylinder / box
    overlaps the oracle's EPA got wrong.  Today's oracle and the GPU agree on each -- the pair's bit on
    the one-warp-per-configuration path, on the tile path and through the signed distance."""
    import json
    from pathlib import Path

    from oracle import oracle
    from ramp_b200 import workloads
    from ramp_b200.desc import ModelDesc

    cases = [json.loads(line) for line in (Path(__file__).parent / "golden" / "sweep_cases.jsonl").read_text().splitlines()]
    assert len(cases) >= 8
    models = {}
    for c in cases:
        key = (c["robot"], c["scene"])
        if key not in models:
            models[key] = _setup(*key)[0]
        rm = models[key]
        d = ModelDesc(rm.model, rm.collision_model)
        q = np.asarray(c["q"], dtype=np.float32)[None]
        ref = oracle.check(d, q[:, : rm.model.nq].astype(np.float64))["pair_dist"][0, c["pair"]]
        assert abs(ref) > BAND, c["note"]
        qd = torch.from_numpy(q).cuda()
        one = bool(rm.gpu.check(qd, pair_mask=True)["pair_mask"][0, c["pair"]])
        filler = torch.from_numpy(workloads.sample_q(rm.model, 9000, seed=5).astype(np.float32)).cuda()
        tile = bool(rm.gpu.check(torch.cat([qd, filler]).contiguous(), pair_mask=True)["pair_mask"][0, c["pair"]])
        dist = float(rm.gpu.distance(qd, [c["pair"]])[0, 0])
        assert one == (ref < 0) and tile == (ref < 0) and (dist < 0) == (ref < 0), (c["note"], ref, one, tile, dist)
        assert abs(dist - ref) < 1e-3 + 0.025 * abs(ref), (c["note"], ref, dist)


def test_scene_update_and_pair_toggle(built):
    """add_object after the handle exists (ramp_model_set_scene) and pair (de)activation."""
    from ramp_b200 import workloads

    rm, rs = _setup("rrr/configs")
    q = workloads.sample_q(rm.model, 4096, seed=11)
    qd = torch.from_numpy(q).cuda()
    before = rm.gpu.check(qd)["hit"].cpu().numpy()
    rs.add_object(workloads.reference_capsule())
    info = rm.gpu.info()
    assert info["n_geoms"] == 6 and info["n_pairs"] == 6 + 5
    after = rm.gpu.check(qd, pair_mask=True)
    assert after["hit"].cpu().numpy().sum() > before.sum()
    mask = after["pair_mask"].cpu().numpy()
    p = int(np.argmax(mask.sum(axis=0)))
    rm.gpu.set_pair_active(p, False)
    off = rm.gpu.check(qd, pair_mask=True)["pair_mask"].cpu().numpy()
    assert not off[:, p].any() and (np.delete(off, p, 1) == np.delete(mask, p, 1)).all()
    rm.gpu.set_pair_active(p, True)


def test_empty_and_errors(built):
    from ramp_b200.exceptions import RampGpuError

    rm, _ = _setup("rrr/configs")
    empty = torch.zeros((0, 4), dtype=torch.float32, device="cuda")
    assert rm.gpu.check(empty)["hit"].numel() == 0
    with pytest.raises(RampGpuError):
        rm.gpu.check(torch.zeros((4, 2), dtype=torch.float32, device="cuda"))  # ld < nq
    with pytest.raises(RampGpuError):
        rm.gpu.forward_kinematics(torch.zeros((4, 3), dtype=torch.float32, device="cuda"), actuated=True)  # no qmap


@pytest.mark.parametrize("name,scene,n", [("rrr/floating_configs", "capsule", 4000), ("rrr_mj/configs", "obstacles", 4000),
                                          ("panda/configs", "capsule", 3000), ("ur5e/configs", "obstacles", 3000)])
def test_signed_distance_matches_oracle(built, name, scene, n):
    """ramp_distance_batch (closed forms, GJK to convergence, EPA) against the oracle's signed
    distance of every pair.  Separations agree to 5e-5 (1 + d) m (FP32 GJK converges to a relative
    1e-5); penetration depths, which both sides reach through an expanding polytope (FP32, 64
    iterations here; curved shapes converge slowly), to 1e-3 m + 2.5 % of the depth (worst seen:
    2.04 % on a cylinder pair of rrr_mj; which sample is worst moves with the last bit of FK)."""
    from oracle import oracle
    from ramp_b200 import workloads
    from ramp_b200.desc import ModelDesc

    rm, rs = _setup(name, scene)
    q = workloads.sample_q(rm.model, n, seed=13)
    got = rm.gpu.distance(torch.from_numpy(q).cuda()).cpu().numpy().astype(np.float64)
    oracle.set_params(1e-8, 128, 1e-8, 255)
    try:
        ref = oracle.check(ModelDesc(rm.model, rm.collision_model), q[:, : rm.model.nq].astype(np.float64))["pair_dist"]
    finally:
        oracle.set_params()
    sep = ref > 1e-4
    pen = ref < -1e-4
    assert sep.sum() > 1000 and pen.sum() > 20
    err_sep = (np.abs(got - ref) / (1.0 + np.abs(ref)))[sep].max()
    err_pen = (np.abs(got - ref) / (1e-3 + 0.025 * np.abs(ref)))[pen].max()
    print(name, dict(pairs=ref.size, separated=int(sep.sum()), penetrating=int(pen.sum()), err_sep=err_sep, err_pen=err_pen))
    assert err_sep < 5e-5 and err_pen < 1.0
    assert ((got <= 0) == (ref <= 0))[sep | pen].all()  # the sign is the collision boolean
    # a sub-selection returns the same columns; RobotState.compute_distance is one of them
    sel = [0, ref.shape[1] - 1]
    sub = rm.gpu.distance(torch.from_numpy(q[:16]).cuda(), sel).cpu().numpy()
    assert np.allclose(sub, got[:16][:, sel], atol=1e-6)


@pytest.mark.parametrize("name,scene,n", [("ur5e/configs", "obstacles", 400000), ("panda/configs", "capsule", 400000),
                                          ("unitree_h1/configs", None, 60000), ("rrr/floating_configs", "capsule", 400000),
                                          ("panda_boxy/configs", "capsule", 400000), ("unitree_h1_boxy/configs", None, 100000)])
def test_large_sweep_flags_and_first_pair(built, name, scene, n):
    """A wide sweep for rare numerical events (FP32 GJK degeneracies show up once in 10^3-10^4
    near-contact pairs): flags and first-pair indices against the oracle, margin band honoured."""
    from oracle import oracle
    from ramp_b200 import workloads
    from ramp_b200.desc import ModelDesc

    rm, _ = _setup(name, scene)
    q = workloads.sample_q(rm.model, n, seed=101)
    res = rm.gpu.check(torch.from_numpy(q).cuda(), first_pair=True)
    hit, first = res["hit"].cpu().numpy(), res["first_pair"].cpu().numpy()
    ref = oracle.check(ModelDesc(rm.model, rm.collision_model), q[:, : rm.model.nq].astype(np.float64))
    pd = ref["pair_dist"]
    outside = np.abs(ref["min_dist"]) > BAND
    clean = (np.abs(pd) > BAND).all(axis=1)
    bad_cfg = int((hit != ref["hit"])[outside].sum())
    bad_first = int((first != ref["first_pair"])[clean].sum())
    print(name, dict(n=n, colliding=int(ref["hit"].sum()), in_band=int((~outside).sum()), bad_cfg=bad_cfg, bad_first=bad_first))
    assert bad_cfg == 0 and bad_first == 0


def test_queue_overflow_is_rechecked(built, monkeypatch, record_property):
    """The GJK work queue is bounded; a tile whose items do not all fit is flagged by the validity
    kernel and recheck_kernel redoes its rows with one warp per configuration (a different GJK
    schedule, so near-contact pairs may flip inside the margin band).  A 100-item queue sends nearly
    every tile down that path -- through ramp_check_batch (ragged batch, flags / first pair / mask)
    and through ramp_check_host (per chunk): the answers still match the oracle outside the band."""
    from oracle import oracle
    from ramp_b200 import workloads
    from ramp_b200.desc import ModelDesc
    from ramp_b200.engine import pinned_empty

    rm, _ = _setup("ur5e/configs", "obstacles")
    q = workloads.sample_q(rm.model, 20011, seed=77)
    monkeypatch.setenv("RAMP_B200_QUEUE_CAP", "100")
    _compare_check(rm, q, lambda **kw: [record_property(k, v) for k, v in kw.items()])
    nq = rm.model.nq
    qh = pinned_empty((len(q), nq), np.float32)
    qh[:] = q[:, :nq]
    hh = pinned_empty((len(q),), np.uint8)
    rm.gpu.check_host_into(qh, hh)
    ref = oracle.check(ModelDesc(rm.model, rm.collision_model), q[:, :nq].astype(np.float64))
    outside = np.abs(ref["min_dist"]) > BAND
    assert (hh.astype(bool) == ref["hit"])[outside].all()


@pytest.mark.gpu
@pytest.mark.parametrize("name,scene", [("ur5e/configs", "obstacles"), ("panda/configs", "capsule"), ("rrr/floating_configs", "capsule")])
def test_unpadded_rows_and_page_locked_host_buffers(built, name, scene):
    """Rows of exactly nq floats (nq not a multiple of 4: a float4 of the tile load straddles two
    rows) give the same flags as padded rows, on the device and through ramp_check_host with
    buffers from ramp_host_alloc; ragged batch sizes exercise the plain-load tail."""
    from ramp_b200 import workloads
    from ramp_b200.engine import pinned_empty

    rm, _ = _setup(name, scene)
    nq = rm.model.nq
    for n in (1, 31, 32, 33, 4097, 300001):
        qp = workloads.sample_q(rm.model, n, seed=5 + n)  # padded rows
        ref = rm.gpu.check(torch.from_numpy(qp).cuda(), first_pair=True)
        tight = np.ascontiguousarray(qp[:, :nq])
        got = rm.gpu.check(torch.from_numpy(tight).cuda(), first_pair=True)
        assert torch.equal(got["hit"], ref["hit"]) and torch.equal(got["first_pair"], ref["first_pair"])
        qh = pinned_empty((n, nq), np.float32)
        qh[:] = tight
        hh = pinned_empty((n,), np.uint8)
        hh[:] = 7
        rm.gpu.check_host_into(qh, hh)
        assert np.array_equal(hh.astype(bool), ref["hit"].cpu().numpy())


@pytest.mark.parametrize("name,scene", [("unitree_h1/configs", None), ("panda/configs", "capsule"), ("ur5e/configs", "obstacles"),
                                        ("rrr/floating_configs", "capsule"), ("rrr/configs", None)])
def test_team_fk_plans_give_the_same_answers(built, monkeypatch, name, scene):
    """Teams of 2 and 4 warps split the kinematic tree between them (FK plans: a run of the joint
    order plus its recomputed ancestors per warp).  Flags, first pairs and pair masks must be
    identical to the one-warp-per-slab kernel, whatever the tree shape."""
    from ramp_b200 import workloads

    rm, _ = _setup(name, scene)
    q = torch.from_numpy(workloads.sample_q(rm.model, 40000, seed=11)).cuda()
    monkeypatch.setenv("RAMP_B200_TEAM", "1")
    ref = rm.gpu.check(q, first_pair=True, pair_mask=True)
    for team in ("2", "4"):
        monkeypatch.setenv("RAMP_B200_TEAM", team)
        got = rm.gpu.check(q, first_pair=True, pair_mask=True)
        for k in ("hit", "first_pair", "pair_mask"):
            assert torch.equal(got[k], ref[k]), (team, k)


def _shape_zoo():
    from ramp_b200.shapes import Box, Capsule, Convex, Cylinder, Ellipsoid, Sphere

    rng = np.random.default_rng(42)
    hull = rng.normal(size=(24, 3)) * np.array([0.12, 0.08, 0.1])
    return [("sphere", Sphere(0.09)), ("capsule", Capsule(0.05, 0.2)), ("box", Box(0.16, 0.1, 0.22)),
            ("cylinder", Cylinder(0.07, 0.18)), ("ellipsoid", Ellipsoid(0.12, 0.07, 0.05)), ("convex", Convex(hull))]


def _shape_zoo_robot():
    """A free-flying body carrying one geometry of each type, in a scene holding one of each."""
    from ramp_b200.kinematics import GeometryObject
    from ramp_b200.robot_model import create_geometry_object, load_robot_model
    from ramp_b200.robot_state import RobotState
    from ramp_b200.se3 import SE3, Quaternion

    rng = np.random.default_rng(7)

    def pose(t):
        v = rng.normal(size=4)
        v /= np.linalg.norm(v)
        return SE3(Quaternion(*v).matrix(), np.asarray(t, dtype=np.float64))

    rm = load_robot_model("zoo::floating")
    cm = rm.collision_model
    n0 = cm.ngeoms
    for k, (name, shape) in enumerate(_shape_zoo()):  # on the free-flyer joint, 0.3 m apart along x
        cm.addGeometryObject(GeometryObject("body_" + name, 1, pose((0.3 * (k - 2.5), 0.0, 0.0)), shape))
    rs = RobotState(rm)
    for k, (name, shape) in enumerate(_shape_zoo()):  # scene: a ring of radius 0.45 m
        a = 2 * np.pi * k / 6
        rs.add_object(create_geometry_object("scene_" + name, shape, pose((0.45 * np.cos(a), 0.45 * np.sin(a), 0.0))))
    assert len(cm.collisionPairs) >= 36 and cm.ngeoms == n0 + 12
    return rm, rs


def test_every_pairing_of_shape_types(built, record_property):
    """every shape RAMP can produce against every other -- sphere, capsule, box,
    cylinder, ellipsoid, convex hull: 36 pairings, random body poses, per-pair decisions against
    the oracle outside the margin band."""
    rm, _ = _shape_zoo_robot()
    q = workloads_sample(rm.model, 60000, seed=3, bound=0.6)
    _compare_check(rm, q, lambda **kw: [record_property(k, v) for k, v in kw.items()])


def test_signed_distance_of_every_pairing_of_shape_types(built):
    """The same scenario through ramp_distance_batch: separations to 5e-5 (1 + d) m, depths to
    1e-3 m + 2.5 % (tolerances of test_signed_distance_matches_oracle), sign = collision flag."""
    from oracle import oracle
    from ramp_b200.desc import ModelDesc

    rm, _ = _shape_zoo_robot()
    q = workloads_sample(rm.model, 20000, seed=4, bound=0.6)
    got = rm.gpu.distance(torch.from_numpy(q).cuda()).cpu().numpy().astype(np.float64)
    oracle.set_params(1e-8, 128, 1e-8, 255)
    try:
        ref = oracle.check(ModelDesc(rm.model, rm.collision_model), q[:, : rm.model.nq].astype(np.float64))["pair_dist"]
    finally:
        oracle.set_params()
    sep, pen = ref > 1e-4, ref < -1e-4
    assert sep.sum() > 1000 and pen.sum() > 200
    err_sep = (np.abs(got - ref) / (1.0 + np.abs(ref)))
    err_pen = (np.abs(got - ref) / (1e-3 + 0.025 * np.abs(ref)))
    worst_sep, worst_pen = err_sep[:, :][sep].max(), err_pen[pen].max()
    per_pair_pen = [float(np.where(pen[:, k], err_pen[:, k], 0).max()) for k in range(ref.shape[1])]
    print(dict(separated=int(sep.sum()), penetrating=int(pen.sum()), err_sep=worst_sep, err_pen=worst_pen,
               worst_pair=int(np.argmax(per_pair_pen))))
    assert worst_sep < 5e-5 and worst_pen < 1.0
    assert ((got <= 0) == (ref <= 0))[sep | pen].all()


def workloads_sample(model, n, seed, bound):
    from ramp_b200 import workloads

    return workloads.sample_q(model, n, seed=seed, position_bound=bound)


def test_host_buffer_call_sections_and_staging(built):
    """ramp_check_host at sizes that span several chunks (2^18 rows) and two sections (2^21 rows:
    one GJK queue and one queue kernel each), with page-locked and with pageable buffers, flags /
    first pair / pair mask: identical to the device-buffer call."""
    from ramp_b200 import workloads
    from ramp_b200.engine import pinned_empty

    rm, _ = _setup("ur5e/configs", "obstacles")
    nq = rm.model.nq
    n = (1 << 21) + 300001
    q = workloads.sample_q(rm.model, n, seed=21)
    ref = rm.gpu.check(torch.from_numpy(q).cuda(), first_pair=True)
    ref_hit, ref_first = ref["hit"].cpu().numpy(), ref["first_pair"].cpu().numpy()
    qh = pinned_empty((n, nq), np.float32)
    qh[:] = q[:, :nq]
    hh = pinned_empty((n,), np.uint8)
    hh[:] = 9
    rm.gpu.check_host_into(qh, hh)
    assert np.array_equal(hh.astype(bool), ref_hit)
    # pageable numpy arrays, all three outputs (a smaller batch: the pair mask is n x pairs)
    m = 600000
    got = rm.gpu.check_host(q[:m, :nq].copy(), first_pair=True, pair_mask=True)
    refm = rm.gpu.check(torch.from_numpy(q[:m]).cuda(), first_pair=True, pair_mask=True)
    assert np.array_equal(got["hit"], ref_hit[:m])
    assert np.array_equal(got["first_pair"], refm["first_pair"].cpu().numpy())
    assert np.array_equal(got["pair_mask"], refm["pair_mask"].cpu().numpy())
    assert np.array_equal(refm["first_pair"].cpu().numpy(), ref_first[:m])


def test_min_distance_output(built):
    """The optional d_min_dist / h_min_dist output of ramp_check_batch / ramp_check_host: smallest
    signed distance over the ACTIVE pairs (a disabled pair does not count), against the oracle."""
    from oracle import oracle
    from ramp_b200 import workloads
    from ramp_b200.desc import ModelDesc

    rm, _ = _setup("ur5e/configs", "obstacles")
    q = workloads.sample_q(rm.model, 3000, seed=31)
    rm.gpu.set_pair_active(3, False)
    try:
        got = rm.gpu.check(torch.from_numpy(q).cuda(), min_dist=True)
        goth = rm.gpu.check_host(q[:, : rm.model.nq].copy(), min_dist=True)
    finally:
        rm.gpu.set_pair_active(3, True)
    pd = oracle.check(ModelDesc(rm.model, rm.collision_model), q[:, : rm.model.nq].astype(np.float64))["pair_dist"]
    ref = np.delete(pd, 3, axis=1).min(axis=1)
    d = got["min_dist"].cpu().numpy().astype(np.float64)
    sep, pen = ref > 1e-4, ref < -1e-4
    assert (np.abs(d - ref) / (1.0 + np.abs(ref)))[sep].max() < 5e-5
    assert (np.abs(d - ref) / (1e-3 + 0.025 * np.abs(ref)))[pen].max() < 1.0
    assert np.array_equal(goth["min_dist"], got["min_dist"].cpu().numpy())
    assert np.array_equal(goth["hit"], got["hit"].cpu().numpy())


@pytest.mark.parametrize("name,scene", [("panda/configs", "capsule"), ("ur5e/configs", "obstacles"),
                                        ("panda_boxy/configs", "capsule"), ("unitree_h1/configs", None), ("rrr_mj/configs", "obstacles")])
def test_pruned_min_distance_equals_the_minimum_of_all_distances(built, name, scene):
    """The d_min_dist output skips GJK / EPA for pairs whose capsule bounds cannot lower the minimum:
    it must be exactly the minimum of the full distance matrix (same arithmetic per pair)."""
    from ramp_b200 import workloads

    rm, _ = _setup(name, scene)
    q = torch.from_numpy(workloads.sample_q(rm.model, 3001, seed=9).astype(np.float32)).cuda()
    full = rm.gpu.distance(q).min(dim=1).values
    got = rm.gpu.check(q, min_dist=True)["min_dist"]
    assert torch.equal(full, got), float((full - got).abs().max())
    assert torch.isfinite(got).all()


@pytest.mark.parametrize("n", [8193 + 17, 20001, 300001])
def test_ragged_batches_never_write_past_the_outputs(built, n):
    """A batch whose size is not a multiple of 32 leaves lanes of the last warp tile without a
    configuration; they replay the tile's first row.  That row is made a colliding one here (its
    hit comes from the GJK stage), so a replayed lane that queued work or stored a result would
    write behind the caller's arrays.  Outputs are views into canary-filled buffers, through
    ramp_check_batch, ramp_check_batch_noalloc and ramp_check_host (with first_pair and mask)."""
    from ramp_b200 import workloads
    from ramp_b200.engine import pinned_empty

    rm, _ = _setup("ur5e/configs", "obstacles")
    gpu = rm.gpu
    npairs = gpu.info()["n_pairs"]
    words = (npairs + 31) // 32
    q = workloads.sample_q(rm.model, n, seed=1000 + n)
    probe = gpu.check(torch.from_numpy(q[:4096]).cuda(), first_pair=True)
    colliding = torch.nonzero(probe["hit"]).flatten().cpu().numpy()
    tail0 = (n // 32) * 32
    q[tail0] = q[colliding[0]]
    qd = torch.from_numpy(q).cuda()
    pad = 64
    stream = torch.cuda.current_stream().cuda_stream

    def canaries():
        hit = torch.full((n + 2 * pad,), 7, dtype=torch.uint8, device="cuda")
        first = torch.full((n + 2 * pad,), 12345, dtype=torch.int32, device="cuda")
        mask = torch.full((n + 2 * pad, words), 0x5A5A5A5A, dtype=torch.int32, device="cuda")
        return hit, first, mask

    def intact(hit, first, mask):
        for t, v in ((hit, 7), (first, 12345), (mask, 0x5A5A5A5A)):
            assert bool((t[:pad] == v).all()) and bool((t[n + pad:] == v).all())

    ref = None
    for entry in ("ramp_check_batch", "ramp_check_batch_noalloc"):
        if entry.endswith("noalloc") and n > 32768:
            continue
        for with_mask in (False, True):
            hit, first, mask = canaries()
            mask[pad : n + pad] = 0
            args = [gpu.handle, qd.data_ptr(), n, qd.shape[1], 2 if with_mask else 0, hit[pad:].data_ptr(), first[pad:].data_ptr()]
            if entry == "ramp_check_batch":
                args.append(None)
            args += [mask[pad:].data_ptr() if with_mask else None, stream]
            assert getattr(gpu.lib, entry)(*args) == 0, gpu.lib.ramp_last_error()
            torch.cuda.synchronize()
            if not with_mask:
                mask[pad : n + pad] = 0x5A5A5A5A
            intact(hit, first, mask)
            assert int(hit[pad + tail0]) == 1
            if ref is None:
                ref = (hit[pad : n + pad].clone(), first[pad : n + pad].clone())
            assert torch.equal(hit[pad : n + pad], ref[0]) and torch.equal(first[pad : n + pad], ref[1])
    # host-buffer call: page-locked arrays with canaries around the live range
    nq = rm.model.nq
    qh = pinned_empty((n, nq), np.float32)
    qh[:] = q[:, :nq]
    hh = pinned_empty((n + 2 * pad,), np.uint8)
    fh = pinned_empty((n + 2 * pad,), np.int32)
    mh = pinned_empty((n + 2 * pad, words), np.uint32)
    hh[:], fh[:], mh[:] = 7, 12345, 0x5A5A5A5A
    rc = gpu.lib.ramp_check_host(gpu.handle, qh.ctypes.data, n, nq, 2, hh[pad:].ctypes.data, fh[pad:].ctypes.data, None,
                                 mh[pad:].ctypes.data)
    assert rc == 0, gpu.lib.ramp_last_error()
    for a, v in ((hh, 7), (fh, 12345), (mh, 0x5A5A5A5A)):
        assert (a[:pad] == v).all() and (a[n + pad:] == v).all()
    assert np.array_equal(hh[pad : n + pad], ref[0].cpu().numpy())
    # first pair with every pair evaluated == first pair of the early-out call (list order)
    assert np.array_equal(fh[pad : n + pad], ref[1].cpu().numpy())


def test_mesh_file_robot_gives_the_box_robot_verdicts(built, tmp_path):
    """URDF <mesh> collision elements (a 12-triangle STL cube and an OBJ cube, scaled) load as
    convex hulls and the device gives the verdicts of the same robot built from <box> elements,
    both matching the oracle outside the margin band."""
    from test_meshes import _write_mesh_robot

    from oracle import oracle
    from ramp_b200 import workloads
    from ramp_b200.desc import ModelDesc
    from ramp_b200.robot_model import load_robot_model
    from ramp_b200.robot_state import RobotState

    prim_path, mesh_path = _write_mesh_robot(tmp_path)
    q = np.random.default_rng(9).uniform(-3, 3, (50000, 2)).astype(np.float32)
    masks = []
    for path in (prim_path, mesh_path):
        rm = load_robot_model(path)
        RobotState(rm).add_object(workloads.reference_capsule())
        res = rm.gpu.check(torch.from_numpy(q).cuda(), pair_mask=True)
        ref = oracle.check(ModelDesc(rm.model, rm.collision_model), q.astype(np.float64))
        out = np.abs(ref["pair_dist"]) > BAND
        mask = res["pair_mask"].cpu().numpy()
        assert ((mask == (ref["pair_dist"] <= 1e-12)) | ~out).all()
        masks.append((mask, out))
    both = masks[0][1] & masks[1][1]
    assert (masks[0][0] == masks[1][0])[both].all()
    assert (masks[0][0].any(axis=0) & ~masks[0][0].all(axis=0)).sum() >= 2


def test_large_batches_with_learnt_queue_size_give_the_same_answers(built):
    """Batches above one slice (2^21 rows): the first call sizes its GJK queue conservatively (3
    items per row, 2^21-row slices); later calls size it from the item counts the earlier ones
    left behind and use larger slices.  Flags and first pairs must not depend on that, and a
    change of scene (other item rates) must not break it."""
    from ramp_b200 import workloads

    rm, rs = _setup("ur5e/configs", "obstacles")
    n = (1 << 21) + 700001
    q = torch.from_numpy(workloads.sample_q(rm.model, n, seed=99)).cuda()
    first = rm.gpu.check(q, first_pair=True)
    torch.cuda.synchronize()
    for _ in range(3):
        again = rm.gpu.check(q, first_pair=True)
        torch.cuda.synchronize()
        assert torch.equal(again["hit"], first["hit"]) and torch.equal(again["first_pair"], first["first_pair"])
    small = rm.gpu.check(q[:300000], first_pair=True)  # one slice: the conservative rule
    assert torch.equal(small["hit"], first["hit"][:300000]) and torch.equal(small["first_pair"], first["first_pair"][:300000])
    # more obstacles -> more work items per row than the estimate knows about: still correct
    rs.add_object(workloads.reference_capsule())
    from ramp_b200.robot_model import create_geometry_object
    from ramp_b200.se3 import SE3
    from ramp_b200.shapes import Box

    rs.add_object(create_geometry_object("wall", Box(0.6, 0.02, 0.8), SE3(np.eye(3), np.asarray([0.3, 0.25, 0.4]))))
    a = rm.gpu.check(q, first_pair=True)
    torch.cuda.synchronize()
    b = rm.gpu.check(q, first_pair=True)
    ref = rm.gpu.check(q[:250000], first_pair=True)
    assert torch.equal(a["hit"], b["hit"]) and torch.equal(a["first_pair"], b["first_pair"])
    assert torch.equal(a["hit"][:250000], ref["hit"]) and torch.equal(a["first_pair"][:250000], ref["first_pair"])
