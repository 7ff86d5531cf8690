"""RobotModel / GroupModel / load_robot_model with the reference's surface
(reference src/ramp/robot_model.py:35-360), on the host containers of this package.

What differs from the reference, on purpose:

* descriptions are read by ramp_b200.description (no Pinocchio/xacrodoc wheels); names
  starting with ``robot-descriptions::`` resolve to the stand-in models of ramp_b200.zoo
  because the upstream assets are fetched from the network by the reference; ``zoo::<name>``
  loads one of the self-contained configurations of ramp_b200.zoo;
* the model owns a lazily built GPU handle (``robot_model.gpu``) that replaces the per-call
  ``createData()`` of reference src/ramp/robot_state.py:376-377;
* q conversion (mimic, continuous) keeps the reference's order, mimic before (cos, sin)
  (reference robot_model.py:406-426), but converts only the slots a setter writes (the
  reference re-applies cos/sin to every continuous joint on every setter).
"""

from __future__ import annotations

import dataclasses
import os
import re
import tomllib
import xml.etree.ElementTree as ET
from dataclasses import dataclass, field
from pathlib import Path

import numpy as np

from ramp_b200 import description, zoo
from ramp_b200.exceptions import (
    MissingAccelerationLimitError,
    MissingBaseLinkError,
    MissingJointError,
    RobotDescriptionNotFoundError,
)
from ramp_b200.kinematics import (
    GeometryModel,
    GeometryObject,
    Model,
    remove_collision_pairs_from_srdf,
)
from ramp_b200.se3 import SE3
from ramp_b200.shapes import CollisionGeometry

ROBOT_DESCRIPTION_PREFIX = "robot-descriptions::"
ZOO_PREFIX = "zoo::"
PINOCCHIO_UNBOUNDED_JOINT = r"JointModelRUB([XYZ])$"
PINOCCHIO_PLANAR_JOINT = "JointModelPlanar"


def get_continuous_joint_indices(model: Model) -> list[int]:
    """q slots holding an angle that Pinocchio stores as (cos, sin)
    (reference src/ramp/pinocchio_utils.py:42-54)."""
    out = []
    for joint in model.joints:
        t = joint.shortname()
        if re.match(PINOCCHIO_UNBOUNDED_JOINT, t):
            out.append(joint.idx_q)
        elif t == PINOCCHIO_PLANAR_JOINT:
            out.append(joint.idx_q + 2)
    return out


def joint_nq(joint) -> int:
    """Number of ACTUATED values of a joint (reference pinocchio_utils.py:67-77)."""
    t = joint.shortname()
    if re.match(PINOCCHIO_UNBOUNDED_JOINT, t):
        return 1
    if t == PINOCCHIO_PLANAR_JOINT:
        return 3
    return joint.nq


def joint_indices(joint) -> list[int]:
    return [joint.idx_q + k for k in range(joint_nq(joint))]


@dataclass(frozen=True, slots=True)
class GroupModel:
    joints: list[str]
    joint_indices: np.ndarray
    joint_position_indices: np.ndarray
    joint_velocity_indices: np.ndarray
    named_states: dict[str, np.ndarray] = field(default_factory=dict)
    tcp_link_name: str | None = None


@dataclass(frozen=True)
class RobotModel:
    model_filename: Path
    base_link: str
    model: Model
    collision_model: GeometryModel
    visual_model: GeometryModel
    groups: dict[str, GroupModel]
    motion_model: dict
    joint_acceleration_limits: dataclasses.InitVar[dict]
    mimic_records: dataclasses.InitVar[list]
    acceleration_limits: np.ndarray = field(init=False)
    joint_names: list[str] = field(init=False)
    mimic_joint_indices: np.ndarray = field(init=False)
    mimic_joint_multipliers: np.ndarray = field(init=False)
    mimic_joint_offsets: np.ndarray = field(init=False)
    mimicked_joint_indices: np.ndarray = field(init=False)
    continuous_joint_indices: np.ndarray = field(init=False)
    _gpu: list = field(init=False, default_factory=list, repr=False, compare=False)

    def __post_init__(self, joint_acceleration_limits, mimic_records):
        st = object.__setattr__
        st(self, "continuous_joint_indices", np.asarray(get_continuous_joint_indices(self.model), dtype=np.int64))
        mimic_ids, mi, mm, mo, msrc = [], [], [], [], []
        for rec in mimic_records or []:
            jid, sid = self.model.getJointId(rec["joint"]), self.model.getJointId(rec["mimicked"])
            a, b = joint_indices(self.model.joints[jid]), joint_indices(self.model.joints[sid])
            assert len(a) == 1, f"Only single joint mimic supported {a}"
            assert len(b) == 1, f"Only single mimicked joint is supported {b}"
            mimic_ids.append(jid)
            mi.append(a[0])
            msrc.append(b[0])
            mm.append(rec.get("multiplier") or 1.0)
            mo.append(rec.get("offset") or 0.0)
        for group in self.groups.values():
            for joint_name in group.joints:
                if self.model.getJointId(joint_name) in mimic_ids:
                    msg = f"Joint '{joint_name}' is a mimic joint, which is not allowed in groups."
                    raise ValueError(msg)
        st(self, "mimic_joint_indices", np.asarray(mi, dtype=np.int64))
        st(self, "mimic_joint_multipliers", np.asarray(mm, dtype=np.float64))
        st(self, "mimic_joint_offsets", np.asarray(mo, dtype=np.float64))
        st(self, "mimicked_joint_indices", np.asarray(msrc, dtype=np.int64))
        names: list[str] = []
        for group in self.groups.values():
            names.extend(group.joints)
        st(self, "joint_names", names)
        acc = []
        if joint_acceleration_limits:
            for joint_name in names:
                if (lim := joint_acceleration_limits.get(joint_name)) is None:
                    raise MissingAccelerationLimitError(names, joint_acceleration_limits.keys())
                acc.append(lim)
        st(self, "acceleration_limits", np.asarray(acc))

    def __getitem__(self, key):
        return self.groups[key]

    def body_geometries(self, name: str):
        fid = self.model.getBodyId(name)
        return [g for g in self.collision_model.geometryObjects if g.parentFrame == fid]

    @property
    def gpu(self):
        """The device-resident model (ramp_b200.engine.GpuModel), built on first use and
        re-synchronised when the collision model was mutated (add_object, pair removal)."""
        from ramp_b200.engine import GpuModel

        if not self._gpu:
            self._gpu.append(GpuModel(self.model, self.collision_model))
        self._gpu[0].sync(self.collision_model)
        return self._gpu[0]


    # ------------------------------------------------------------------ batched entry points
    # (SURVEY section 8b "added": the calls a user of the reference makes once per state, for a
    # whole batch; q rows are Pinocchio configurations, float32 [N, nq] on the host or the device)
    def check_collisions(self, q):
        """bool[N]: any enabled pair in collision (batched RobotState.check_collision)."""
        return self.gpu.check(q)["hit"]

    def forward_kinematics(self, q):
        """float32[N, J, 3, 4]: oMi of every movable joint (batched pinocchio.forwardKinematics)."""
        return self.gpu.forward_kinematics(q)

    def check_edges(self, group_name: str, q0, q1, reference_state=None):
        """(valid bool[E], first_invalid int32[E]): OMPL's DiscreteMotionValidator over E edges
        between ACTUATED states of ``group_name``; q slots outside the group come from
        ``reference_state`` (default: the neutral configuration)."""
        from ramp_b200 import spaces

        ref = None if reference_state is None else reference_state.qpos
        self.gpu.set_qmap(spaces.make_qmap(self, group_name, ref))
        res = self.gpu.check_edges(q0, q1, spaces.make_space(self, group_name))
        return res["valid"], res["first_invalid"]


    def ik_params(self, group_name: str, **kw):
        """RampIkParams for the group's tool frame: its parent joint and placement, the joint
        limits of the group's actuated values (continuous joints are unbounded), and the
        reference's MAX_ITERATIONS / MAX_ERROR (src/ramp/constants.py:10-11) unless overridden."""
        from ramp_b200.desc import IkParams

        group = self.groups[group_name]
        assert group.tcp_link_name, f"tcp_link_name is not defined for group '{group_name}'"
        frame = self.model.frames[self.model.getFrameId(group.tcp_link_name)]
        cont = set(int(c) for c in self.continuous_joint_indices)
        lo = [-np.inf if int(i) in cont else self.model.lowerPositionLimit[int(i)] for i in group.joint_position_indices]
        hi = [np.inf if int(i) in cont else self.model.upperPositionLimit[int(i)] for i in group.joint_position_indices]
        return IkParams(frame.parentJoint - 1, frame.placement.rt34().reshape(-1), lo, hi, **kw)

    def differential_ik_batch(self, group_name: str, seeds, target_poses, reference_state=None, **kw) -> dict:
        """Batched differential IK: every row of ``seeds`` (actuated values of the
        group) is iterated towards ``target_poses`` ([n, 3, 4] / [3, 4] world poses of the group's
        tool frame).  Returns ``q``, ``err``, ``ok``, ``iters`` (torch tensors on the device).  The
        goal-state generator of reference examples/so_arm100_motion_planning.py:50-54 in one call."""
        from ramp_b200 import spaces

        ref = None if reference_state is None else reference_state.qpos
        self.gpu.set_qmap(spaces.make_qmap(self, group_name, ref))
        return self.gpu.ik_batch(seeds, target_poses, self.ik_params(group_name, **kw))


def make_groups(model: Model, configs: dict) -> dict[str, GroupModel]:
    """reference src/ramp/robot_model.py:187-246."""
    groups = {}
    link_names = [f.name for f in model.frames]
    for group_name, gc in configs["group"].items():
        joints = gc["joints"]
        tcp = gc.get("tcp_link_name")
        if tcp is not None:
            assert tcp in link_names, f"Group {group_name} TCP link '{tcp}' not in link names: {link_names}"
        for joint_name in joints:
            if joint_name not in model.names:
                msg = f"Joint '{joint_name}' for group '{group_name}' not in model joints: {list(model.names)}"
                raise MissingJointError(msg)
        idx, qidx, vidx = [], [], []
        for joint_name in joints:
            j = model.joints[model.getJointId(joint_name)]
            idx.append(j.id)
            qidx.extend(joint_indices(j))
            vidx.append(j.idx_v)
        named = {}
        for state_name, values in gc.get("named_states", {}).items():
            assert len(values) == len(qidx), (
                f"Named state '{state_name}' has {len(values)} joint positions, expected {len(qidx)} for {joints}")
            named[state_name] = np.asarray(values, dtype=np.float64)
        groups[group_name] = GroupModel(joints, np.asarray(idx), np.asarray(qidx), np.asarray(vidx), named, tcp)
    return groups


def _finish(model_filename, models, configs, disabled) -> RobotModel:
    model, collision_model, mimics = models
    collision_model.addAllCollisionPairs()
    if disabled:
        remove_collision_pairs_from_srdf(model, collision_model, disabled)
    link_names = [f.name for f in model.frames]
    base_link = configs["robot"]["base_link"]
    if base_link not in link_names:
        msg = f"Base link '{base_link}' not in link names: {link_names}"
        raise MissingBaseLinkError(msg)
    return RobotModel(Path(model_filename), base_link, model, collision_model, GeometryModel(),
                      make_groups(model, configs), configs.get("motion_model", {}),
                      configs.get("acceleration_limits", {}), mimics)


def _bare_configs(model: Model, motion_models, acceleration_limits) -> dict:
    return {"robot": {"base_link": "universe"},
            "group": {"default": {"joints": [n for n in model.names if n != "universe"]}},
            "acceleration_limits": acceleration_limits or {}, "motion_model": motion_models or {}}


def _resolve_description(desc: str, base: Path):
    """-> (filename, (Model, GeometryModel, mimics)) for a TOML ``description`` value."""
    if desc.startswith(ROBOT_DESCRIPTION_PREFIX):
        name = desc[len(ROBOT_DESCRIPTION_PREFIX):]
        models = zoo.standin(name)
        if models is None:
            raise RobotDescriptionNotFoundError(name)
        return Path(f"standin::{name}.urdf"), models
    return base / desc, None


def load_robot_model(config_path, motion_models: dict | None = None,
                     acceleration_limits: dict | None = None) -> RobotModel:
    """Load from ``zoo::<name>``, a TOML config, or a bare URDF / xacro / MJCF file
    (reference src/ramp/robot_model.py:249-310)."""
    s = str(config_path)
    if s.startswith(ZOO_PREFIX):
        name = s[len(ZOO_PREFIX):]
        if name not in zoo.CONFIGS:
            raise FileNotFoundError(f"No zoo configuration '{name}'; known: {sorted(zoo.CONFIGS)}")
        cfg = zoo.CONFIGS[name]
        if "standin" in cfg:
            models = zoo.standin(cfg["standin"])
            filename = Path(f"standin::{cfg['standin']}.urdf")
        elif cfg["kind"] == "urdf":
            models = description.build_models_from_urdf(description.render_xacro(cfg["text"]))
            filename = Path(f"zoo::{name}.urdf")
        else:
            models = description.build_models_from_mjcf(ET.fromstring(cfg["text"]))
            filename = Path(f"zoo::{name}.xml")
        configs = _bare_configs(models[0], motion_models, acceleration_limits) if cfg.get("bare") else cfg
        return _finish(filename, models, configs, cfg["disabled"])
    if s.startswith(ROBOT_DESCRIPTION_PREFIX):
        filename, models = _resolve_description(s, Path("."))
        return _finish(filename, models, _bare_configs(models[0], motion_models, acceleration_limits), [])
    config_path = Path(config_path)
    if not config_path.exists():
        msg = f"File does not exist: {config_path}"
        raise FileNotFoundError(msg)
    if config_path.suffix == ".toml":
        with open(config_path, "rb") as f:
            configs = tomllib.load(f)
        mappings = configs["robot"].get("mappings", {})
        filename, models = _resolve_description(configs["robot"]["description"], config_path.parent)
        if models is None:
            models = description.load_description(filename, mappings)
        disabled = []
        if srdf := configs["robot"].get("disable_collisions"):
            disabled = description.read_disabled_collisions(
                description.render_xacro(config_path.parent / srdf, mappings))
        if os.environ.get("LOG_LEVEL", "INFO").upper() == "DEBUG":
            print(configs)
        return _finish(filename, models, configs, disabled)
    if config_path.suffix in (".urdf", ".xacro", ".xml"):
        models = description.load_description(config_path, {})
        return _finish(config_path, models, _bare_configs(models[0], motion_models, acceleration_limits), [])
    msg = f"Unknown robot description file type: {config_path}"
    raise ValueError(msg)


# ------------------------------------------------------------- q conversions
def apply_pinocchio_mimic_joints(robot_model: RobotModel, q: np.ndarray):
    """reference src/ramp/robot_model.py:419-426."""
    if len(robot_model.mimic_joint_indices) == 0:
        return
    q[robot_model.mimic_joint_indices] = (
        q[robot_model.mimicked_joint_indices] * robot_model.mimic_joint_multipliers + robot_model.mimic_joint_offsets)


def apply_pinocchio_continuous_joints(robot_model: RobotModel, q: np.ndarray, slots=None):
    """Angle -> (cos, sin) for the given continuous slots (default: all, as the reference does,
    src/ramp/robot_model.py:406-416)."""
    idx = robot_model.continuous_joint_indices if slots is None else np.asarray(slots, dtype=np.int64)
    if len(idx) == 0:
        return
    q[idx], q[idx + 1] = np.cos(q[idx]), np.sin(q[idx])


def from_pinocchio_qpos_continuous(robot_model: RobotModel, q):
    """(cos, sin) -> angle in place (reference src/ramp/robot_model.py:378-388)."""
    idx = robot_model.continuous_joint_indices
    if len(idx) == 0:
        return
    q[idx] = np.arctan2(q[idx + 1], q[idx])


def get_converted_qpos(robot_model: RobotModel, qpos: np.ndarray) -> np.ndarray:
    qpos = np.copy(qpos)
    from_pinocchio_qpos_continuous(robot_model, qpos)
    return qpos


def set_positions(robot_model: RobotModel, q: np.ndarray, position_indices, values) -> None:
    """Write actuated values into a Pinocchio q: scatter, mimic, then (cos, sin) for the
    continuous slots among the written ones (and mimic targets)."""
    position_indices = np.asarray(position_indices, dtype=np.int64)
    q[position_indices] = values
    apply_pinocchio_mimic_joints(robot_model, q)
    written = set(position_indices.tolist()) | set(robot_model.mimic_joint_indices.tolist())
    slots = [int(c) for c in robot_model.continuous_joint_indices if int(c) in written]
    apply_pinocchio_continuous_joints(robot_model, q, slots)


def as_pinocchio_qpos(robot_model: RobotModel, reference_qpos, group_name: str, group_qpos) -> np.ndarray:
    """reference src/ramp/robot_model.py:363-375."""
    q = np.copy(reference_qpos)
    set_positions(robot_model, q, robot_model[group_name].joint_position_indices, group_qpos)
    return q


def create_geometry_object(name: str, geometry: CollisionGeometry, pose: SE3) -> GeometryObject:
    """A universe-attached scene object (reference src/ramp/robot_model.py:438-455)."""
    obj = GeometryObject(name, 0, pose, geometry)
    obj.meshColor = np.array([0.0, 1.0, 0.0, 1.0])
    return obj
