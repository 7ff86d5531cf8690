"""MotionPlanner over the batched GPU validator: the reference's planning known answers
(reference tests/motion_planner_test.py:17-65 -- "a plan is found") plus checks the reference
cannot make: every waypoint and every edge of the returned path is valid under the CPU oracle."""

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def _oracle_path_valid(rm, planner, path):
    from oracle import oracle
    from ramp_b200 import spaces
    from ramp_b200.desc import ModelDesc

    d = ModelDesc(rm.model, rm.collision_model)
    q = np.stack([s.qpos for s in path])
    res = oracle.check(d, q)
    states = np.stack([planner.as_state(s) for s in path])
    nd = [spaces.valid_segment_count(planner.space, a, b) for a, b in zip(states[:-1], states[1:])]
    return res, max(nd)


@pytest.mark.parametrize("cfg,scene,goal,timeout", [
    ("rrr/configs", None, [0.0, 1.0, 1.0], 2.0),
    ("rrr/configs", "capsule", [0.0, 2.0, 1.4], 5.0),
    ("rrr/planar_configs", "capsule", [1.0, -0.5, 1.57, 0.5, 0.25, 0.1], 5.0),
    ("rrr/floating_configs", "capsule", [1.0, 0.5, 1.0, 1.0, 0.0, 0.0, 0.0, 0.5, 0.25, 0.1], 5.0),
    ("panda/configs", "capsule", [0.0, 0.25, 0.0, -1.25, 0.0, 1.5, 0.785], 5.0),
])
def test_plan_found_and_valid_under_oracle(built, cfg, scene, goal, timeout):
    from ramp_b200 import MotionPlanner, RobotState, load_robot_model, workloads

    rm = load_robot_model("zoo::" + cfg)
    start = RobotState.from_named_state(rm, "arm", "home")
    if scene:
        start.add_object(workloads.reference_capsule())
    planner = MotionPlanner(rm, "arm")
    path = planner.plan(start, goal, timeout=timeout)
    assert path is not None, "Expected a plan to be found"
    assert np.allclose(planner.as_state(path[0]), planner.as_state(start), atol=1e-6)
    res, max_nd = _oracle_path_valid(rm, planner, path)
    # interpolated path: consecutive states at most one valid segment apart
    assert max_nd <= 1
    # every waypoint collision free (or inside the +-1e-4 margin band)
    assert (~res["hit"] | (np.abs(res["min_dist"]) <= 1e-4)).all()
    goal_state = start.clone()
    goal_state.set_group_qpos("arm", goal)
    assert np.allclose(path[-1].qpos, goal_state.qpos, atol=1e-5)


def test_invalid_start_or_goal_returns_none(built):
    from ramp_b200 import MotionPlanner, RobotState, load_robot_model

    rm = load_robot_model("zoo::rrr/configs")
    planner = MotionPlanner(rm, "arm")
    start = RobotState.from_named_state(rm, "arm", "home")
    assert planner.plan(start, [0.0, 2.9, 2.9]) is None       # goal folds the arm into itself
    assert planner.plan(start, [0.0, 4.0, 0.0]) is None       # goal out of bounds
    assert planner.is_state_valid(start)
    bad = start.clone()
    bad.set_group_qpos("arm", [0.0, 2.9, 2.9])
    assert not planner.is_state_valid(bad)


def test_check_motion_overloads_follow_ompl(built):
    from oracle import oracle
    from ramp_b200 import MotionPlanner, RobotState, spaces, workloads
    from ramp_b200.desc import ModelDesc

    rm, rs = workloads.build("rrr_float")
    planner = MotionPlanner(rm, "arm")
    planner.set_reference_state(rs)
    d = ModelDesc(rm.model, rm.collision_model)
    rng = np.random.default_rng(5)
    q0 = spaces.sample_uniform(planner.space, 30, rng)
    q1 = spaces.sample_uniform(planner.space, 30, rng)
    q1[:, :3] = q0[:, :3] + rng.normal(scale=0.3, size=(30, 3))
    ref = oracle.check_edges(d, planner._qmap, planner.space, q0, q1)
    seen_invalid = 0
    for e in range(30):
        lv = []
        ok = planner.validator.checkMotion(q0[e], q1[e], lv)
        assert ok == planner.validator.checkMotion(q0[e], q1[e])
        if ok != bool(ref["valid"][e]):
            continue  # a margin-band state; covered by test_edges_match_oracle's accounting
        if not ok and ref["nd"][e] > 0:
            seen_invalid += 1
            assert lv[1] == pytest.approx((ref["first_invalid"][e] - 1) / ref["nd"][e], abs=1.5 / ref["nd"][e])
            assert np.allclose(lv[0], spaces.interpolate(planner.space, q0[e], q1[e], lv[1]))
    assert seen_invalid >= 3


def test_get_frame_pose_and_collision_pairs(built):
    """RobotState.get_frame_pose (FK + frame placement) and the all-pairs report."""
    from ramp_b200 import RobotState, load_robot_model
    from ramp_b200.compute_disable_collisions import default_collisions, get_collision_pairs

    rm = load_robot_model("zoo::rrr/configs")
    rs = RobotState.from_named_state(rm, "arm", "home")
    rs.set_group_qpos("arm", [0.3, 0.5, -0.2])
    pose = rs.get_frame_pose("end_effector")
    # joint3 origin + 0.3 along link3's z (pitch 0.3 total, yaw 0.3)
    j3 = np.array([0.4 * np.sin(0.5) * np.cos(0.3), 0.4 * np.sin(0.5) * np.sin(0.3), 0.525 + 0.4 * np.cos(0.5)])
    z = np.array([np.sin(0.3) * np.cos(0.3), np.sin(0.3) * np.sin(0.3), np.cos(0.3)])
    assert np.allclose(pose.translation, j3 + 0.3 * z, atol=1e-5)
    assert pose.np.shape == (4, 4)
    rs.set_group_qpos("arm", [0.0, 2.9, 2.9])
    pairs = get_collision_pairs(rm, rs.qpos)
    assert ("link1", "link3") in pairs or ("link1", "end_effector") in pairs
    assert default_collisions(rm) == []  # nothing collides at the neutral configuration
    assert rs.check_collision() and rs.check_collision(verbose=True)


def test_sampled_always_never_pairs(built):
    """ALWAYS / NEVER classification (the reference's TODO): rrr_mj has no SRDF, so its adjacent
    link boxes overlap in every configuration; far-apart pairs never touch."""
    from ramp_b200 import load_robot_model
    from ramp_b200.compute_disable_collisions import DisabledReason, disable_collision, sampled_collisions

    rm = load_robot_model("zoo::rrr_mj/configs")
    found = sampled_collisions(rm, n_samples=1 << 15)
    always = {(a, b) for a, b, r in found if r == DisabledReason.ALWAYS}
    assert {("link1", "link2"), ("link2", "link3")} <= always
    n0 = len(rm.collision_model.collisionPairs)
    disable_collision(rm, found)
    assert len(rm.collision_model.collisionPairs) == n0 - len(found)
    # with those pairs gone the home state of the MJCF robot is valid (as its URDF twin's is)
    from ramp_b200 import RobotState

    assert not RobotState.from_named_state(rm, "arm", "home").check_collision()


def test_projection_batch_matches_frame_pose(built):
    from ramp_b200 import MotionPlanner, RobotState, load_robot_model

    rm = load_robot_model("zoo::panda/configs")
    planner = MotionPlanner(rm, "arm")
    rs = RobotState.from_named_state(rm, "arm", "home")
    planner.set_reference_state(rs)
    rng = np.random.default_rng(1)
    states = rng.uniform(-1.5, 1.5, (64, 7))
    states[:, 3] = rng.uniform(-3.0, -0.1, 64)
    got = planner.project(states)
    for i in (0, 17, 63):
        rs.set_group_qpos("arm", states[i])
        assert np.allclose(got[i], rs.get_frame_pose("panda_hand_tcp").translation, atol=2e-5)


def test_path_clearance_matches_oracle(built):
    """clearance along a planned path (state cost of the reference's
    PathClearanceObjective) = smallest signed pair distance per interpolated state, one batched
    call; against the oracle's min_dist."""
    from ramp_b200 import MotionPlanner, RobotState, load_robot_model, workloads

    rm = load_robot_model("zoo::rrr/configs")
    start = RobotState.from_named_state(rm, "arm", "home")
    start.add_object(workloads.reference_capsule())
    planner = MotionPlanner(rm, "arm")
    path = planner.plan(start, [0.0, 2.0, 1.4], timeout=5.0)
    assert path is not None
    states = [planner.as_state(s) for s in path]
    clr = planner.path_clearance(states)
    res, _ = _oracle_path_valid(rm, planner, path)
    assert len(clr) >= len(path) and clr.min() > -1e-4
    # the planned path is already interpolated, so the states coincide with the waypoints
    assert np.allclose(np.sort(clr)[:3], np.sort(res["min_dist"])[:3], atol=5e-5)


def test_robot_model_batch_entry_points(built):
    """RobotModel.check_collisions / forward_kinematics / check_edges (SURVEY 8b "added") agree
    with the per-state reference-shaped calls."""
    from ramp_b200 import MotionPlanner, RobotState, load_robot_model, workloads

    rm = load_robot_model("zoo::rrr/configs")
    start = RobotState.from_named_state(rm, "arm", "home")
    start.add_object(workloads.reference_capsule())
    q = workloads.sample_q(rm.model, 64, seed=2)
    hit = rm.check_collisions(q).cpu().numpy()
    for i in range(0, 64, 8):
        s = start.clone()
        s.qpos = q[i, : rm.model.nq].astype(np.float64)
        assert s.check_collision() == bool(hit[i])
    assert tuple(rm.forward_kinematics(q).shape) == (64, rm.model.njoints - 1, 3, 4)
    planner = MotionPlanner(rm, "arm")
    planner.set_reference_state(start)
    rng = np.random.default_rng(0)
    a, b = rng.uniform(-1.5, 1.5, (32, 3)), rng.uniform(-1.5, 1.5, (32, 3))
    valid, first = rm.check_edges("arm", a, b, start)
    for i in range(0, 32, 4):
        assert planner.validator.checkMotion(a[i], b[i]) == bool(valid[i])


def test_are_states_valid_matches_is_state_valid(built):
    """MotionPlanner.are_states_valid (vectorised bounds + one batched check) equals
    is_state_valid state by state (reference src/ramp/motion_planner.py:612-625)."""
    from ramp_b200 import MotionPlanner, RobotState, load_robot_model, spaces, workloads

    rm = load_robot_model("zoo::rrr/configs")
    start = RobotState.from_named_state(rm, "arm", "home")
    start.add_object(workloads.reference_capsule())
    planner = MotionPlanner(rm, "arm")
    planner.set_reference_state(start)
    rng = np.random.default_rng(3)
    states = spaces.sample_uniform(planner.space, 64, rng)
    states[::9] *= 3.0  # some out of bounds
    got = planner.are_states_valid(states)
    for i in range(0, 64, 3):
        s = start.clone()
        s.set_group_qpos("arm", states[i])
        assert planner.is_state_valid(s) == bool(got[i]), i
    assert got.any() and not got.all()


@pytest.mark.parametrize("cfg,scene", [("rrr_mj/configs", None), ("panda/configs", "capsule"), ("ur5e/configs", "obstacles")])
def test_sampled_always_never_against_the_oracle(built, cfg, scene):
    """ALWAYS = a pair collides in EVERY sampled configuration, NEVER = in none
    (the reference's TODOs, src/ramp/compute_disable_collisions.py:229-230).  The GPU's per-pair
    AND / OR reduction against the oracle's ``pair_dist <= 0`` over the SAME samples; a pair whose
    classification hinges on a sample inside the +-1e-4 m band is exempt."""
    from oracle import oracle
    from ramp_b200 import load_robot_model, workloads
    from ramp_b200.compute_disable_collisions import DisabledReason, get_collision_pair_names, sampled_collisions
    from ramp_b200.desc import ModelDesc
    from ramp_b200.robot_state import RobotState

    rm = load_robot_model("zoo::" + cfg)
    rs = RobotState(rm)
    if scene == "capsule":
        rs.add_object(workloads.reference_capsule())
    elif scene == "obstacles":
        for o in workloads.obstacle_scene():
            rs.add_object(o)
    n, seed = 1 << 14, 5
    found = sampled_collisions(rm, n_samples=n, seed=seed, batch=n)
    q = workloads.sample_q_torch(rm.model, n, seed=seed, device=rm.gpu.device).cpu().numpy()  # the samples the tool drew
    pd = oracle.check(ModelDesc(rm.model, rm.collision_model), q[:, : rm.model.nq].astype(np.float64))["pair_dist"]
    hit, band = pd <= 1e-12, np.abs(pd) <= 1e-4
    cm = rm.collision_model
    got = {(a, b): r for a, b, r in found}
    checked = 0
    for k, pair in enumerate(cm.collisionPairs):
        names = get_collision_pair_names(rm.model, cm, pair)
        want = DisabledReason.ALWAYS if hit[:, k].all() else (DisabledReason.NEVER if not hit[:, k].any() else None)
        # could a flipped in-band sample change the class?  then either answer is acceptable
        firm_hit, firm_free = (hit[:, k] & ~band[:, k]), (~hit[:, k] & ~band[:, k])
        ambiguous = band[:, k].any() and not (firm_hit.any() and firm_free.any())
        if not ambiguous:
            assert got.get(names) == want, (names, got.get(names), want)
            checked += 1
    assert checked >= len(cm.collisionPairs) // 2  # (rrr_mj: three adjacent box pairs touch exactly in every pose)
    assert any(r == DisabledReason.NEVER for r in got.values()) or any(r == DisabledReason.ALWAYS for r in got.values())


@pytest.mark.parametrize("cfg", ["panda/configs", "rrr/floating_configs", "rrr/planar_configs"])
def test_projection_matches_oracle_fk(built, cfg):
    """the batched projection (TCP position of every state, one ramp_fk_batch call)
    against the oracle: oMi[parent joint of the TCP frame] * frame.placement, reference
    src/ramp/motion_planner.py:485-506 / src/ramp/robot_state.py:194-212."""
    from oracle import oracle
    from ramp_b200 import MotionPlanner, RobotState, load_robot_model, spaces
    from ramp_b200.desc import ModelDesc

    rm = load_robot_model("zoo::" + cfg)
    planner = MotionPlanner(rm, "arm")
    rs = RobotState.from_named_state(rm, "arm", "home")
    planner.set_reference_state(rs)
    states = spaces.sample_uniform(planner.space, 500, np.random.default_rng(2))
    got = planner.project(states)
    q = oracle.qmap_apply(planner._qmap, states.astype(np.float32).astype(np.float64))
    omi = oracle.fk(ModelDesc(rm.model, rm.collision_model), q)
    frame = rm.model.frames[rm.model.getFrameId(rm["arm"].tcp_link_name)]
    j = frame.parentJoint - 1
    want = omi[:, j, :, :3] @ frame.placement.translation + omi[:, j, :, 3]
    assert got.shape == want.shape and np.abs(got - want).max() < 1e-5
