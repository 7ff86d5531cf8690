"""The description readers: (i) the zoo's generated text gives the expected tables, (ii) when
the reference tree is present (this container; never on the GPU box) its own robot files load
to tables IDENTICAL to the zoo's, which pins the zoo against the reference's data."""

from pathlib import Path

import numpy as np
import pytest

from ramp_b200 import description
from ramp_b200.compute_disable_collisions import DisabledReason, adjacent_collisions
from ramp_b200.desc import ModelDesc
from ramp_b200.exceptions import MissingAccelerationLimitError, MissingBaseLinkError, MissingJointError, RobotDescriptionNotFoundError
from ramp_b200.robot_model import load_robot_model

REF = Path("/root/reference")
needs_ref = pytest.mark.skipif(not (REF / "robots").exists(), reason="reference tree not present")


def _tables(rm):
    d = ModelDesc(rm.model, rm.collision_model)
    joints = [(j.type, j.parent, j.idx_q, tuple(j.axis), tuple(np.round(list(j.placement), 12))) for j in d.joints[: d.struct.n_joints]]
    geoms = [(g.type, g.parent_joint, tuple(g.params), tuple(np.round(list(g.placement), 12))) for g in d.geoms[: d.n_geoms]]
    names = [g.name for g in rm.collision_model.geometryObjects]
    frames = [(f.name, f.parentJoint, f.type) for f in rm.model.frames]
    return joints, geoms, d.pairs.tolist(), names, frames, rm.model.names


@needs_ref
@pytest.mark.parametrize("zoo_name,ref_path", [
    ("rrr/configs", "robots/rrr/configs.toml"),
    ("rrr/planar_configs", "robots/rrr/planar_configs.toml"),
    ("rrr/floating_configs", "robots/rrr/floating_configs.toml"),
    ("planar_rrr/configs", "robots/planar_rrr/configs.toml"),
    ("rrr_mj/configs", "robots/rrr_mj/configs.toml"),
    ("acrobot/configs", "robots/acrobot/configs.toml"),
    ("double_pendulum", "tests/double_pendulum.toml"),
    ("floating", "robots/floating.urdf"),
    ("planar", "robots/planar.urdf"),
])
def test_reference_files_load_to_the_zoo_tables(zoo_name, ref_path):
    a = load_robot_model("zoo::" + zoo_name)
    b = load_robot_model(REF / ref_path)
    assert _tables(a) == _tables(b)
    assert set(a.groups) == set(b.groups)
    for g in a.groups:
        assert a[g].joints == b[g].joints and a[g].tcp_link_name == b[g].tcp_link_name
        assert list(a[g].joint_position_indices) == list(b[g].joint_position_indices)
        assert {k: v.tolist() for k, v in a[g].named_states.items()} == {k: v.tolist() for k, v in b[g].named_states.items()}
    assert np.array_equal(a.mimic_joint_indices, b.mimic_joint_indices)
    assert np.array_equal(a.continuous_joint_indices, b.continuous_joint_indices)


@needs_ref
@pytest.mark.parametrize("zoo_name,ref_path", [("panda/configs", "robots/panda/configs.toml"),
                                               ("ur5e/configs", "robots/ur5e/configs.toml"),
                                               ("unitree_h1/configs", "robots/unitree_h1/configs.toml")])
def test_reference_tomls_resolve_to_standins(zoo_name, ref_path):
    """robot-descriptions:: names resolve to the stand-ins; the reference's SRDF (rendered by
    the xacro subset, macros included) removes exactly the pairs the zoo's list removes."""
    a = load_robot_model("zoo::" + zoo_name)
    b = load_robot_model(REF / ref_path)
    assert _tables(a)[2] == _tables(b)[2]
    assert {g: a[g].joints for g in a.groups} == {g: b[g].joints for g in b.groups}


def test_adjacent_pairs_match_reference_test():
    # reference tests/disable_collisions_test.py:12-41
    def adj(name):
        rm = load_robot_model("zoo::" + name)
        return set(adjacent_collisions(rm))

    A = DisabledReason.ADJACENT
    assert adj("rrr/configs") == {("base_link", "link1", A), ("link1", "link2", A), ("link2", "link3", A),
                                  ("link3", "end_effector", A)}
    assert adj("acrobot/configs") == {("upper_link", "lower_link", A)}
    assert adj("planar_rrr/configs") == {("base_link", "link1", A), ("link1", "link2", A), ("link2", "link3", A)}


def test_joint_models_follow_pinocchio_rules():
    rm = load_robot_model("zoo::double_pendulum")
    kinds = [j.shortname() for j in rm.model.joints[1:]]
    # continuous +x -> RUBX; axis (0,-1,0) is not a positive unit axis -> RevoluteUnaligned
    assert kinds == ["JointModelRUBX", "JointModelRUBX", "JointModelRevoluteUnaligned"]
    assert rm.model.nq == 5 and list(rm.continuous_joint_indices) == [0, 2]
    rm = load_robot_model("zoo::rrr/floating_configs")
    assert rm.model.joints[1].shortname() == "JointModelFreeFlyer" and rm.model.nq == 10
    # fixed joints fold links into the parent joint: link3 and end_effector share joint3
    g = {o.name: o.parentJoint for o in rm.collision_model.geometryObjects}
    assert g["link3_0"] == g["end_effector_0"] and g["mobile_base_0"] == g["base_link_0"] == 1
    rm = load_robot_model("zoo::rrr_mj/configs")
    # MJCF sizes are half extents: cylinder size="0.1 0.025" -> r 0.1, half length 0.025
    base = rm.collision_model.geometryObjects[0].geometry
    assert base.radius == pytest.approx(0.1) and base.halfLength == pytest.approx(0.025)
    assert rm.model.lowerPositionLimit[0] == pytest.approx(-3.14)


def test_xacro_subset():
    xml = """<robot xmlns:xacro="http://www.ros.org/wiki/xacro" name="t">
      <xacro:arg name="p" default="a_"/><xacro:property name="l" value="2.0"/>
      <xacro:macro name="m"><disable_collisions link1="$(arg p)x" link2="$(arg p)y"/></xacro:macro>
      <xacro:if value="${l > 1}"><link name="$(arg p)big"><collision><geometry><box size="1 1 ${l/2}"/></geometry></collision></link></xacro:if>
      <xacro:unless value="${l > 1}"><link name="small"/></xacro:unless>
      <xacro:m/></robot>"""
    root = description.render_xacro(xml, {"p": "b_"})
    assert [l.get("name") for l in root.findall("link")] == ["b_big"]
    assert root.find("link/collision/geometry/box").get("size") == "1 1 1.0"
    assert description.read_disabled_collisions(root) == [("b_x", "b_y")]


def test_loader_errors(tmp_path):
    # reference tests/robot_test.py:85-114
    with pytest.raises(FileNotFoundError):
        load_robot_model(tmp_path / "missing.toml")
    with pytest.raises(RobotDescriptionNotFoundError):
        load_robot_model("robot-descriptions::non_existing_description")
    from ramp_b200 import zoo

    base = zoo.CONFIGS["rrr/configs"]
    for key, patch, exc in (("_e1", {"robot": {"base_link": "nope"}}, MissingBaseLinkError),
                            ("_e2", {"group": {"arm": {"joints": ["joint1", "jointX"]}}}, MissingJointError),
                            ("_e3", {"acceleration_limits": {"joint1": 1.0}}, MissingAccelerationLimitError)):
        zoo.CONFIGS[key] = {**base, **patch}
        try:
            with pytest.raises(exc):
                load_robot_model("zoo::" + key)
        finally:
            del zoo.CONFIGS[key]
