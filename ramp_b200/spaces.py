"""Group -> actuated layout, q map and OMPL-equivalent state space description.

The reference builds one OMPL subspace per group joint (reference
src/ramp/motion_planner.py:264-359) and moves values between OMPL states and Pinocchio q
through ``from_ompl_state`` (:89-162) and ``set_group_qpos`` (robot_state.py:109-119).  The
batched path needs the same facts as flat tables: which actuated column feeds which q slot,
which slots are angles stored as (cos, sin), the mimic couplings, and per-subspace type,
bounds and extent.
"""

from __future__ import annotations

import re

import numpy as np

from ramp_b200 import desc as D
from ramp_b200.kinematics import neutral

REVOLUTE = r"JointModelR([XYZ])$"
PRISMATIC = r"JointModelP([XYZ])$"
UNBOUNDED = r"JointModelRUB([XYZ])$"


def make_qmap(robot_model, group_name: str | None, reference_qpos: np.ndarray | None = None) -> D.QMap:
    """Actuated columns of ``group_name`` (None: all groups in order, as
    ``RobotState.set_actuated_qpos`` does, reference robot_state.py:131-140) -> Pinocchio q.

    ``reference_qpos`` supplies every slot the group does not write (already in Pinocchio
    form).  Continuous conversion is applied to the written continuous slots and to mimic
    targets only (the reference's re-application to other groups is a bug).
    """
    model = robot_model.model
    ref = neutral(model) if reference_qpos is None else np.asarray(reference_qpos, dtype=np.float64)
    groups = [robot_model[group_name]] if group_name else list(robot_model.groups.values())
    slots = np.concatenate([np.asarray(g.joint_position_indices, dtype=np.int64) for g in groups])
    written = set(slots.tolist()) | set(robot_model.mimic_joint_indices.tolist())
    cont = [int(c) for c in robot_model.continuous_joint_indices if int(c) in written]
    return D.QMap(model.nq, slots, ref, robot_model.mimic_joint_indices, robot_model.mimicked_joint_indices,
                  robot_model.mimic_joint_multipliers, robot_model.mimic_joint_offsets, cont)


def make_space(robot_model, group_name: str, *, position_bound: float = 10.0, fraction: float = 0.01,
               check_bounds: bool = True) -> D.SpaceDesc:
    """The compound space ``MotionPlanner.__init__`` builds, as a flat description.

    Revolute / prismatic / revolute-unaligned joints: RealVector(1) with the model's limits;
    unbounded revolute: SO2; planar: SE2 with x, y in +-10; free-flyer: SE3 with xyz in +-10
    (reference motion_planner.py:268-359).  Dubins / Reeds-Shepp / Vana / Owen motion models
    are not discretised here (their interpolation stays with OMPL on the host).
    """
    model = robot_model.model
    subs, col = [], 0
    for idx in robot_model[group_name].joint_indices:
        joint = model.joints[int(idx)]
        name = model.names[int(idx)]
        t = joint.shortname()
        if robot_model.motion_model.get(name):
            raise ValueError(f"motion model '{robot_model.motion_model[name]}' of joint '{name}' is not supported "
                             "by the batched validator")
        if re.match(REVOLUTE, t) or re.match(PRISMATIC, t) or t == "JointModelRevoluteUnaligned":
            subs.append({"type": D.SPACE_REAL, "col": col, "dim": 1,
                         "lo": [model.lowerPositionLimit[joint.idx_q]], "hi": [model.upperPositionLimit[joint.idx_q]]})
            col += 1
        elif re.match(UNBOUNDED, t):
            subs.append({"type": D.SPACE_SO2, "col": col})
            col += 1
        elif t == "JointModelPlanar":
            subs.append({"type": D.SPACE_SE2, "col": col, "lo": [-position_bound] * 2, "hi": [position_bound] * 2})
            col += 3
        elif t == "JointModelFreeFlyer":
            subs.append({"type": D.SPACE_SE3, "col": col, "lo": [-position_bound] * 3, "hi": [position_bound] * 3})
            col += 7
        elif t == "JointModelSpherical":
            subs.append({"type": D.SPACE_SO3, "col": col})
            col += 4
        elif t == "JointModelTranslation":
            subs.append({"type": D.SPACE_REAL, "col": col, "dim": 3, "lo": [-np.inf] * 3, "hi": [np.inf] * 3})
            col += 3
        else:
            msg = f"Unknown joint type: '{t}' for joint '{name}'"
            raise ValueError(msg)
    return D.SpaceDesc(subs, fraction, check_bounds)


def sample_uniform(space: D.SpaceDesc, n: int, rng: np.random.Generator, position_bound: float = 1.0) -> np.ndarray:
    """Uniform samples of the space in actuated-column form (test / benchmark input)."""
    out = np.zeros((n, space.n_cols))
    for s in space.subspaces:
        c = s["col"]
        t = s["type"]
        if t == D.SPACE_REAL:
            for k in range(s["dim"]):
                lo = s["lo"][k] if np.isfinite(s["lo"][k]) else -position_bound
                hi = s["hi"][k] if np.isfinite(s["hi"][k]) else position_bound
                out[:, c + k] = rng.uniform(lo, hi, n)
        elif t == D.SPACE_SO2:
            out[:, c] = rng.uniform(-np.pi, np.pi, n)
        elif t == D.SPACE_SE2:
            out[:, c : c + 2] = rng.uniform(-position_bound, position_bound, (n, 2))
            out[:, c + 2] = rng.uniform(-np.pi, np.pi, n)
        elif t in (D.SPACE_SE3, D.SPACE_SO3):
            o = 3 if t == D.SPACE_SE3 else 0
            if o:
                out[:, c : c + 3] = rng.uniform(-position_bound, position_bound, (n, 3))
            v = rng.normal(size=(n, 4))
            out[:, c + o : c + o + 4] = v / np.linalg.norm(v, axis=1, keepdims=True)
    return out
