"""Known-answer facts the reference's own tests and examples pin for this path,
checked against the CPU oracle and the host model layer.  No GPU needed."""

import numpy as np
import pytest

from oracle import oracle
from ramp_b200 import workloads
from ramp_b200.desc import ModelDesc
from ramp_b200.kinematics import GeometryObject, neutral
from ramp_b200.robot_model import as_pinocchio_qpos, load_robot_model
from ramp_b200.robot_state import RobotState
from ramp_b200.se3 import SE3, Quaternion
from ramp_b200.shapes import Sphere


def _valid(rm, qs):
    d = ModelDesc(rm.model, rm.collision_model)
    return ~oracle.check(d, np.asarray(qs, dtype=np.float64))["hit"]


def test_rrr_pair_count_and_home_goal_valid():
    # reference tests/motion_planner_test.py:19-23, robots/rrr/rrr.srdf.xacro:8-11
    rm = load_robot_model("zoo::rrr/configs")
    assert rm.collision_model.ngeoms == 5 and len(rm.collision_model.collisionPairs) == 6
    home = RobotState.from_named_state(rm, "arm", "home").qpos
    goal = as_pinocchio_qpos(rm, home, "arm", [0.0, 1.0, 1.0])
    assert _valid(rm, [home, goal]).all()


@pytest.mark.parametrize("cfg,goal", [
    ("rrr/planar_configs", [1.0, -0.5, 1.57, 0.5, 0.25, 0.1]),            # motion_planner_test.py:28-44
    ("rrr/floating_configs", [1.0, 0.5, 1.0, 1.0, 0.0, 0.0, 0.0, 0.5, 0.25, 0.1]),  # :47-65
    ("rrr/configs", [0.0, 2.0, 1.4]),                                     # examples/motion_planning.py:28
])
def test_start_and_goal_valid_with_reference_capsule(cfg, goal):
    rm = load_robot_model("zoo::" + cfg)
    rs = RobotState.from_named_state(rm, "arm", "home")
    n_robot = rm.collision_model.ngeoms
    n_pairs = len(rm.collision_model.collisionPairs)
    rs.add_object(workloads.reference_capsule())
    # add_object pairs the object with every robot geometry (robot_state.py:322-331)
    assert len(rm.collision_model.collisionPairs) == n_pairs + n_robot
    g = as_pinocchio_qpos(rm, rs.qpos, "arm", goal)
    assert _valid(rm, [rs.qpos, g]).all()


def test_capsule_quaternion_is_used_unnormalised():
    # Quaternion(0.707, 0.707, 0, 0) is not unit; Eigen's toRotationMatrix does not normalise
    r = Quaternion(0.707, 0.707, 0.0, 0.0).matrix()
    assert abs(np.linalg.det(r) - 1.0) > 1e-4
    assert np.allclose(r[:, 0], [1.0, 0.0, 0.0], atol=1e-3)
    assert np.allclose(r @ [0, 0, 1], [0, -1, 0], atol=1e-3)  # the capsule axis lies along -y


def test_airplane_lattice_scene():
    # reference examples/airplane_motion_planning.py:26-51: 27 spheres, box airplane
    rm = load_robot_model("zoo::floating")
    start = RobotState.from_actuated_qpos(rm, [0.0, 0.1, 0.0, 0.0, 0.0, 0.0, 1.0])
    for i in range(3):
        for j in range(3):
            for k in range(3):
                start.add_object(GeometryObject(f"sphere_{i}x{j}x{k}", 0,
                                                SE3(Quaternion.Identity(), np.asarray([i * 0.25, j * 0.25, k * 0.25])),
                                                Sphere(0.05)))
    cm = rm.collision_model
    assert cm.ngeoms == 28 and len(cm.collisionPairs) == 27  # never scene-scene pairs
    rot = Quaternion(0.707, 0.707, 0.0, 0.0).normalize()
    goal = RobotState.from_actuated_qpos(rm, [0.4, 0.4, 0.4, rot.x, rot.y, rot.z, rot.w])
    # the start pose (0, 0.1, 0) overlaps the sphere at the origin? box half extents
    # (0.05, 0.0375, 0.025), sphere r = 0.05 at distance 0.1 - 0.0375 = 0.0625 > 0.05: free
    v = _valid(rm, [start.qpos, goal.qpos])
    assert v[0] and v[1]
    d = ModelDesc(rm.model, cm)
    md = oracle.check(d, start.qpos[None])["min_dist"][0]
    assert md == pytest.approx(0.1 - 0.0375 - 0.05, abs=1e-9)


def test_double_pendulum_mimic_and_continuous_bookkeeping():
    # reference tests/robot_state_test.py:41-57
    rm = load_robot_model("zoo::double_pendulum")
    rs = RobotState(rm)
    assert np.allclose(rs.actuated_qpos(), [0.0, 0.0])
    rs.set_group_qpos("arm", [0.2, 0.1])
    assert np.allclose(rs.actuated_qpos(), [0.2, 0.1])
    assert np.allclose(rs.qpos, RobotState.from_actuated_qpos(rm, rs.actuated_qpos()).qpos)
    rs2 = rs.clone()
    rs2.set_group_qpos("arm", [0.5, 0.1])
    assert not np.allclose(rs2.actuated_qpos(), rs.actuated_qpos())
    assert np.allclose(rs2.actuated_qpos(), rs2.group_qpos("arm"))
    assert np.allclose(rs2.qpos[rm.mimic_joint_indices], [-0.05])
    # Pinocchio layout: (cos, sin) pairs for the two continuous joints
    assert np.allclose(rs2.qpos[:4], [np.cos(0.5), np.sin(0.5), np.cos(0.1), np.sin(0.1)])


def test_mimic_joint_in_group_rejected():
    # reference tests/mimic_test.py:13-19
    from ramp_b200 import zoo

    cfg = dict(zoo.CONFIGS["double_pendulum"])
    cfg["group"] = {"arm": {"joints": ["joint1", "joint3"]}}
    zoo.CONFIGS["_bad_mimic"] = cfg
    try:
        with pytest.raises(ValueError, match="mimic joint"):
            load_robot_model("zoo::_bad_mimic")
    finally:
        del zoo.CONFIGS["_bad_mimic"]


def test_panda_standin_facts():
    # reference tests/mimic_test.py:22-43 and SURVEY §8a pair arithmetic
    rm = load_robot_model("zoo::panda/configs")
    assert len(rm.mimic_joint_indices) == 1
    m = rm.model
    assert m.names[m.getJointId("panda_finger_joint2")] == "panda_finger_joint2"
    assert rm.mimicked_joint_indices[0] == m.joints[m.getJointId("panda_finger_joint1")].idx_q
    assert m.nq == 9 and rm.collision_model.ngeoms == 11 and len(rm.collision_model.collisionPairs) == 21
    rs = RobotState.from_named_state(rm, "arm", "home")
    rs.add_object(workloads.reference_capsule())
    assert _valid(rm, [rs.qpos]).all()  # the example plans from home with the capsule present


def test_h1_group_bookkeeping():
    # reference tests/robot_state_test.py:18-38 (stand-in model, same group layout)
    rm = load_robot_model("zoo::unitree_h1/configs")
    rs = RobotState.from_named_state(rm, "left_arm", "home")
    assert np.allclose(rs.actuated_qpos(), np.zeros(8))
    rs.set_group_qpos("left_arm", [0.5] * 4)
    rs.set_group_qpos("right_arm", [-0.5] * 4)
    assert np.allclose(rs.actuated_qpos(), [0.5] * 4 + [-0.5] * 4)
    rs3 = rs.clone()
    rs3.set_group_qpos("left_arm", [0.0, 0.0, 0.0, 0.2])
    assert np.allclose(rs3.qpos, RobotState.from_actuated_qpos(rm, rs3.actuated_qpos()).qpos)
    assert rm.model.nq == 26 and len(rm.collision_model.collisionPairs) == 190 - 13


def test_fk_hand_computed_rrr():
    rm = load_robot_model("zoo::rrr/configs")
    d = ModelDesc(rm.model, rm.collision_model)
    q = np.array([[0.3, 0.5, -0.2]])
    omi = oracle.fk(d, q)[0]
    # joint3 origin: 0.025 + 0.5 up, then 0.4 along link2's z tilted by Ry(0.5), yawed by Rz(0.3)
    p = np.array([0.4 * np.sin(0.5) * np.cos(0.3), 0.4 * np.sin(0.5) * np.sin(0.3), 0.525 + 0.4 * np.cos(0.5)])
    assert np.allclose(omi[2, :, 3], p, atol=1e-12)
    # total pitch 0.5 - 0.2 about the yawed y axis
    z_axis = omi[2, :, 2]
    assert np.allclose(z_axis, [np.sin(0.3) * np.cos(0.3), np.sin(0.3) * np.sin(0.3), np.cos(0.3)], atol=1e-12)


def test_neutral_and_first_pair_order():
    rm = load_robot_model("zoo::rrr/configs")
    d = ModelDesc(rm.model, rm.collision_model)
    # fold the arm onto itself: link3 / end effector reach back into link1 and the base
    q = np.array([[0.0, 2.9, 2.9]])
    r = oracle.check(d, q)
    assert r["hit"][0]
    hits = np.flatnonzero(r["pair_dist"][0] <= 1e-12)
    assert r["first_pair"][0] == hits[0]  # collisionPairIndex = first colliding pair in list order
    assert (neutral(rm.model) == 0).all()
