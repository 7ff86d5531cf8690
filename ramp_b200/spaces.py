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


# ----------------------------------------------------------------------------------------
# Host-side (numpy, FP64) statement of the OMPL state-space maths the planner loop needs
# between validator calls: distance, interpolation, bounds.  The batched
# validator itself runs these in FP32 on the device (csrc/edges.cuh).
def _so2_dist(a: float, b: float) -> float:
    d = abs(a - b)
    return 2 * np.pi - d if d > np.pi else d


def _so2_interp(a: float, b: float, t: float) -> float:
    diff = b - a
    if abs(diff) <= np.pi:
        return a + diff * t
    diff = 2 * np.pi - diff if diff > 0 else -2 * np.pi - diff
    v = a - diff * t
    if v > np.pi:
        v -= 2 * np.pi
    elif v < -np.pi:
        v += 2 * np.pi
    return v


def _so3_arc(a, b) -> float:
    dq = abs(float(np.dot(a, b)))
    return 0.0 if dq > 1.0 - 1e-9 else float(np.arccos(dq))


def _so3_slerp(a, b, t: float):
    th = _so3_arc(a, b)
    if th <= np.finfo(float).eps:
        return np.array(a, dtype=np.float64)
    s0, s1 = np.sin((1 - t) * th), np.sin(t * th)
    if np.dot(a, b) < 0:
        s1 = -s1
    return (np.asarray(a) * s0 + np.asarray(b) * s1) / np.sin(th)


def _parts(s):
    """(n_real, rotation kind, rotation offset) of a subspace."""
    t = s["type"]
    if t == D.SPACE_REAL:
        return s["dim"], None, 0
    if t == D.SPACE_SO2:
        return 0, "so2", 0
    if t == D.SPACE_SE2:
        return 2, "so2", 2
    if t == D.SPACE_SE3:
        return 3, "so3", 3
    return 0, "so3", 0


def distance(space: D.SpaceDesc, a, b) -> float:
    """CompoundStateSpace::distance with RAMP's weights (1 per joint subspace; SE2 = R2 + 0.5 SO2,
    SE3 = R3 + SO3)."""
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    total = 0.0
    for s in space.subspaces:
        c = s["col"]
        nreal, rot, off = _parts(s)
        if nreal:
            total += float(np.linalg.norm(a[c : c + nreal] - b[c : c + nreal]))
        if rot == "so2":
            total += (0.5 if s["type"] == D.SPACE_SE2 else 1.0) * _so2_dist(a[c + off], b[c + off])
        elif rot == "so3":
            total += _so3_arc(a[c + off : c + off + 4], b[c + off : c + off + 4])
    return total


def distance_many(space: D.SpaceDesc, many, b) -> np.ndarray:
    """``distance(space, many[i], b)`` for every row of ``many`` (a planner's nearest-neighbour query).
    Spaces made of one-dimensional real subspaces only (revolute / prismatic joints: the sum of
    the joint differences) are done in one numpy expression, the others row by row."""
    many = np.asarray(many, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    subs = [_parts(s) for s in space.subspaces]
    if all(nreal == 1 and rot is None for nreal, rot, _ in subs):
        cols = [s["col"] for s in space.subspaces]
        return np.abs(many[:, cols] - b[cols]).sum(axis=1)
    return np.array([distance(space, a, b) for a in many])


def interpolate(space: D.SpaceDesc, a, b, t: float) -> np.ndarray:
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    out = np.array(a, dtype=np.float64)
    for s in space.subspaces:
        c = s["col"]
        nreal, rot, off = _parts(s)
        out[c : c + nreal] = a[c : c + nreal] + (b[c : c + nreal] - a[c : c + nreal]) * t
        if rot == "so2":
            out[c + off] = _so2_interp(a[c + off], b[c + off], t)
        elif rot == "so3":
            out[c + off : c + off + 4] = _so3_slerp(a[c + off : c + off + 4], b[c + off : c + off + 4], t)
    return out


def valid_segment_count(space: D.SpaceDesc, a, b) -> int:
    """max over (sub)components of ceil(distance / (fraction * extent))."""
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    f = space.struct.longest_valid_segment_fraction
    nd = 0
    for s in space.subspaces:
        c = s["col"]
        nreal, rot, off = _parts(s)
        if nreal:
            ext = float(np.linalg.norm(np.asarray(s["hi"][:nreal]) - np.asarray(s["lo"][:nreal])))
            if ext > 0 and np.isfinite(ext):
                nd = max(nd, int(np.ceil(np.linalg.norm(a[c : c + nreal] - b[c : c + nreal]) / (f * ext))))
        if rot == "so2":
            nd = max(nd, int(np.ceil(_so2_dist(a[c + off], b[c + off]) / (f * np.pi))))
        elif rot == "so3":
            nd = max(nd, int(np.ceil(_so3_arc(a[c + off : c + off + 4], b[c + off : c + off + 4]) / (f * 0.5 * np.pi))))
    return nd


def satisfies_bounds(space: D.SpaceDesc, a) -> bool:
    a = np.asarray(a, dtype=np.float64)
    eps = np.finfo(float).eps
    for s in space.subspaces:
        c = s["col"]
        nreal, rot, off = _parts(s)
        for k in range(nreal):
            if a[c + k] - eps > s["hi"][k] or a[c + k] + eps < s["lo"][k]:
                return False
        if rot == "so2" and not (-np.pi - eps <= a[c + off] <= np.pi + eps):
            return False
    return True


def satisfies_bounds_batch(space: D.SpaceDesc, states) -> np.ndarray:
    """``satisfies_bounds`` for every row of ``states`` [N, n_cols] at once (bool[N])."""
    a = np.atleast_2d(np.asarray(states, dtype=np.float64))
    eps = np.finfo(float).eps
    ok = np.ones(len(a), dtype=bool)
    for s in space.subspaces:
        c = s["col"]
        nreal, rot, off = _parts(s)
        for k in range(nreal):
            ok &= (a[:, c + k] - eps <= s["hi"][k]) & (a[:, c + k] + eps >= s["lo"][k])
        if rot == "so2":
            ok &= (a[:, c + off] >= -np.pi - eps) & (a[:, c + off] <= np.pi + eps)
    return ok


def maximum_extent(space: D.SpaceDesc) -> float:
    ext = 0.0
    for s in space.subspaces:
        nreal, rot, _ = _parts(s)
        if nreal:
            e = float(np.linalg.norm(np.asarray(s["hi"][:nreal]) - np.asarray(s["lo"][:nreal])))
            ext += e if np.isfinite(e) else 0.0
        if rot == "so2":
            ext += (0.5 if s["type"] == D.SPACE_SE2 else 1.0) * np.pi
        elif rot == "so3":
            ext += 0.5 * np.pi
    return ext
