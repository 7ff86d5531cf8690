"""OMPL's discretisation rule in the oracle and in the host-side space
maths: segment counts, interpolation, bounds, and the equivalence between OMPL's two
``checkMotion`` overloads and the "smallest invalid index" the batched validator returns."""

from collections import deque

import numpy as np
import pytest

from oracle import oracle
from ramp_b200 import spaces, workloads
from ramp_b200.desc import SPACE_REAL, SPACE_SE2, SPACE_SE3, SPACE_SO2, ModelDesc, SpaceDesc
from ramp_b200.robot_model import load_robot_model
from ramp_b200.robot_state import RobotState


def test_real_vector_rule():
    sp = SpaceDesc([{"type": SPACE_REAL, "col": 0, "dim": 1, "lo": [-3.14159], "hi": [3.14159]},
                    {"type": SPACE_REAL, "col": 1, "dim": 1, "lo": [0.0], "hi": [0.04]}])
    # nd = max_j ceil(|dq_j| / (0.01 * range_j))
    assert oracle.edge_nd(sp, [0.0, 0.0], [1.0, 0.0]) == int(np.ceil(1.0 / (0.01 * 6.28318)))
    assert oracle.edge_nd(sp, [0.0, 0.0], [0.01, 0.02]) == 50
    assert oracle.edge_nd(sp, [0.3, 0.01], [0.3, 0.01]) == 0
    assert np.allclose(oracle.edge_interpolate(sp, [0.0, 0.0], [1.0, 0.04], 0.25), [0.25, 0.01])


def test_so2_wraps_the_short_way():
    sp = SpaceDesc([{"type": SPACE_SO2, "col": 0}])
    assert oracle.edge_nd(sp, [3.0], [-3.0]) == int(np.ceil((2 * np.pi - 6.0) / (0.01 * np.pi)))
    mid = oracle.edge_interpolate(sp, [3.0], [-3.0], 0.5)[0]
    assert abs(abs(mid) - np.pi) < 1e-12
    assert oracle.edge_interpolate(sp, [3.0], [-3.0], 0.25)[0] == pytest.approx(3.0 + 0.25 * (2 * np.pi - 6.0))


def test_se3_extents():
    sp = SpaceDesc([{"type": SPACE_SE3, "col": 0, "lo": [-10.0] * 3, "hi": [10.0] * 3}])
    a = np.array([0, 0, 0, 0, 0, 0, 1.0])
    b = np.array([1.0, 0, 0, 0, 0, 0, 1.0])
    assert oracle.edge_nd(sp, a, b) == int(np.ceil(1.0 / (0.01 * np.sqrt(1200.0))))  # lvs = 0.3464
    c = np.array([0, 0, 0, np.sin(0.25), 0, 0, np.cos(0.25)])  # rotation by 0.5 rad: arc 0.25
    assert oracle.edge_nd(sp, a, c) == int(np.ceil(0.25 / (0.01 * 0.5 * np.pi)))
    q = oracle.edge_interpolate(sp, a, c, 0.5)
    assert np.allclose(q[3:], [np.sin(0.125), 0, 0, np.cos(0.125)])
    # antipodal quaternion = same rotation: zero arc
    assert oracle.edge_nd(sp, a, np.array([0, 0, 0, 0, 0, 0, -1.0])) == 0


def test_se2_and_bounds():
    sp = SpaceDesc([{"type": SPACE_SE2, "col": 0, "lo": [-10.0] * 2, "hi": [10.0] * 2}])
    assert oracle.edge_nd(sp, [0, 0, 0], [1, 0, 0]) == int(np.ceil(1.0 / (0.01 * np.sqrt(800.0))))
    assert oracle.edge_nd(sp, [0, 0, 0], [0, 0, 1.0]) == int(np.ceil(1.0 / (0.01 * np.pi)))
    assert oracle.satisfies_bounds(sp, [10.0, -10.0, 3.14])
    assert not oracle.satisfies_bounds(sp, [10.1, 0.0, 0.0])
    assert not oracle.satisfies_bounds(sp, [0.0, 0.0, 3.2])


@pytest.mark.parametrize("cfg", ["rrr/configs", "rrr/planar_configs", "rrr/floating_configs", "planar_rrr/configs",
                                 "panda/configs"])
def test_host_space_maths_equal_the_oracle(cfg):
    rm = load_robot_model("zoo::" + cfg)
    sp = spaces.make_space(rm, "arm")
    rng = np.random.default_rng(0)
    a, b = spaces.sample_uniform(sp, 200, rng), spaces.sample_uniform(sp, 200, rng)
    for x, y in zip(a, b):
        assert spaces.valid_segment_count(sp, x, y) == oracle.edge_nd(sp, x, y)
        t = rng.uniform()
        assert np.allclose(spaces.interpolate(sp, x, y, t), oracle.edge_interpolate(sp, x, y, t), atol=1e-12)
        assert spaces.satisfies_bounds(sp, x) == oracle.satisfies_bounds(sp, x)


def _ompl_check_motion_bool(valid, nd):
    """DiscreteMotionValidator::checkMotion(s1, s2): s2 first, then interior states in
    breadth-first bisection order.  valid(j) is the validity of state j/nd (j = nd: s2)."""
    if not valid(nd):
        return False
    if nd >= 2:
        pos = deque([(1, nd - 1)])
        while pos:
            lo, hi = pos.popleft()
            mid = (lo + hi) // 2
            if not valid(mid):
                return False
            if lo < mid:
                pos.append((lo, mid - 1))
            if hi > mid:
                pos.append((mid + 1, hi))
    return True


def _ompl_check_motion_last_valid(valid, nd):
    """The pair overload: linear order, returns (ok, last valid parameter)."""
    for j in range(1, nd):
        if not valid(j):
            return False, (j - 1) / nd
    if not valid(nd):
        return False, (nd - 1) / nd if nd > 0 else 0.0
    return True, None


def test_first_invalid_index_implies_both_ompl_overloads():
    rm, rs = workloads.build("rrr_float")
    sp = spaces.make_space(rm, "arm")
    qmap = spaces.make_qmap(rm, "arm", RobotState(rm).qpos)
    d = ModelDesc(rm.model, rm.collision_model)
    rng = np.random.default_rng(3)
    q0 = spaces.sample_uniform(sp, 40, rng)
    q1 = spaces.sample_uniform(sp, 40, rng)
    q1[:, :3] = q0[:, :3] + rng.normal(scale=0.3, size=(40, 3))
    res = oracle.check_edges(d, qmap, sp, q0, q1)
    assert (~res["valid"]).sum() > 3 and res["valid"].sum() > 3
    for e in range(40):
        nd = int(res["nd"][e])

        def valid(j, e=e, nd=nd):
            st = q1[e] if j >= nd else oracle.edge_interpolate(sp, q0[e], q1[e], j / nd)
            if not oracle.satisfies_bounds(sp, st):
                return False
            return not oracle.check(d, oracle.qmap_apply(qmap, st))["hit"][0]

        first = int(res["first_invalid"][e])
        assert _ompl_check_motion_bool(valid, nd) == (first < 0) == bool(res["valid"][e])
        ok, t = _ompl_check_motion_last_valid(valid, nd)
        assert ok == (first < 0)
        if not ok and nd > 0:
            assert t == pytest.approx((first - 1) / nd)


def test_batched_bounds_equal_the_per_state_rule():
    """spaces.satisfies_bounds_batch is the vectorised form of satisfies_bounds (StateSpace::
    satisfiesBounds per subspace), including values exactly on a bound."""
    from ramp_b200 import spaces
    from ramp_b200.robot_model import load_robot_model

    rng = np.random.default_rng(5)
    for cfg, group in (("rrr/configs", "arm"), ("rrr/planar_configs", "arm"), ("rrr/floating_configs", "arm")):
        rm = load_robot_model("zoo::" + cfg)
        space = spaces.make_space(rm, group)
        a = spaces.sample_uniform(space, 400, rng)
        a[::7] *= 1.5  # push some rows out of bounds
        for s in space.subspaces:  # and some exactly onto a bound
            if len(s["hi"]) and np.isfinite(s["hi"][0]):
                a[3, s["col"]] = s["hi"][0]
        want = np.array([spaces.satisfies_bounds(space, row) for row in a])
        assert np.array_equal(spaces.satisfies_bounds_batch(space, a), want)
        assert want.any() and not want.all()


def test_distance_many_matches_distance():
    """The planner's vectorised nearest-neighbour distances equal CompoundStateSpace::distance row by
    row: a space of 1-D real subspaces (fast path) and one with SO2 / SE3 parts (row-by-row path)."""
    from ramp_b200 import load_robot_model, spaces

    rng = np.random.default_rng(5)
    for cfg in ("fr3/configs", "rrr/planar_configs", "rrr/floating_configs"):
        rm = load_robot_model("zoo::" + cfg)
        sp = spaces.make_space(rm, "arm")
        pts = spaces.sample_uniform(sp, 40, rng)
        got = spaces.distance_many(sp, pts[1:], pts[0])
        ref = np.array([spaces.distance(sp, p, pts[0]) for p in pts[1:]])
        assert np.allclose(got, ref, rtol=0, atol=1e-12)
