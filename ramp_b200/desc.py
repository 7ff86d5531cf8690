"""ctypes mirror of include/ramp_b200.h and the builders that flatten a host model into it.

SURVEY §7 step 1: a ``Model`` + ``GeometryModel`` become POD tables -- joints in Pinocchio
order (parents precede children), geometries with their parent joint and placement, the
pair list in ``collisionPairs`` order, the convex vertices, the actuated->q map
(reference src/ramp/robot_model.py:406-426) and the OMPL space description
(reference src/ramp/motion_planner.py:264-359).  The same descriptors feed the CUDA library
and, in tests only, the CPU oracle.
"""

from __future__ import annotations

import ctypes as C
import re

import numpy as np

from ramp_b200.kinematics import GeometryModel, Model

ABI_VERSION = 1

JOINT_TYPE = {
    "JointModelRX": 1, "JointModelRY": 2, "JointModelRZ": 3, "JointModelRevoluteUnaligned": 4,
    "JointModelRUBX": 5, "JointModelRUBY": 6, "JointModelRUBZ": 7,
    "JointModelRevoluteUnboundedUnaligned": 8,
    "JointModelPX": 9, "JointModelPY": 10, "JointModelPZ": 11, "JointModelPrismaticUnaligned": 12,
    "JointModelPlanar": 13, "JointModelFreeFlyer": 14, "JointModelSpherical": 15,
    "JointModelSphericalZYX": 16, "JointModelTranslation": 17,
}  # fmt: skip

SPACE_REAL, SPACE_SO2, SPACE_SE2, SPACE_SE3, SPACE_SO3 = range(5)

Q_ACTUATED = 1
ALL_PAIRS = 2


class RampJoint(C.Structure):
    _fields_ = [("type", C.c_int32), ("parent", C.c_int32), ("idx_q", C.c_int32),
                ("reserved", C.c_int32), ("axis", C.c_double * 3), ("placement", C.c_double * 12)]


class RampGeom(C.Structure):
    _fields_ = [("type", C.c_int32), ("parent_joint", C.c_int32), ("vert_offset", C.c_int32),
                ("vert_count", C.c_int32), ("params", C.c_double * 4),
                ("placement", C.c_double * 12)]


class RampQMap(C.Structure):
    _fields_ = [("n_act", C.c_int32), ("act_slot", C.POINTER(C.c_int32)),
                ("n_mimic", C.c_int32), ("mimic_dst", C.POINTER(C.c_int32)),
                ("mimic_src", C.POINTER(C.c_int32)), ("mimic_mult", C.POINTER(C.c_double)),
                ("mimic_off", C.POINTER(C.c_double)),
                ("n_cont", C.c_int32), ("cont_slot", C.POINTER(C.c_int32)),
                ("ref_q", C.POINTER(C.c_double))]


class RampModelDesc(C.Structure):
    _fields_ = [("abi_version", C.c_int32), ("nq", C.c_int32),
                ("n_joints", C.c_int32), ("joints", C.POINTER(RampJoint)),
                ("n_geoms", C.c_int32), ("geoms", C.POINTER(RampGeom)),
                ("n_pairs", C.c_int32), ("pairs", C.POINTER(C.c_int32)),
                ("n_verts", C.c_int32), ("verts", C.POINTER(C.c_double))]


class RampSubspace(C.Structure):
    _fields_ = [("type", C.c_int32), ("col", C.c_int32), ("dim", C.c_int32),
                ("reserved", C.c_int32), ("lo", C.c_double * 3), ("hi", C.c_double * 3)]


class RampSpaceDesc(C.Structure):
    _fields_ = [("n_sub", C.c_int32), ("sub", C.POINTER(RampSubspace)),
                ("longest_valid_segment_fraction", C.c_double),
                ("check_bounds", C.c_int32), ("reserved", C.c_int32)]


class RampIkParams(C.Structure):
    _fields_ = [("frame_parent_joint", C.c_int32), ("max_iterations", C.c_int32), ("frame_placement", C.c_double * 12),
                ("damping", C.c_double), ("step", C.c_double), ("tolerance", C.c_double), ("max_step", C.c_double),
                ("lower", C.POINTER(C.c_double)), ("upper", C.POINTER(C.c_double))]


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _ptr(a: np.ndarray, ctype):
    return a.ctypes.data_as(C.POINTER(ctype))


def fill_geoms(objs, vert_base: int = 0):
    """GeometryObjects -> (RampGeom array, vertex array [V,3]).  ``parentJoint`` is a
    Pinocchio joint id (0 = universe); the tables index movable joints from 0."""
    arr = (RampGeom * max(len(objs), 1))()
    verts = []
    nv = vert_base
    for i, g in enumerate(objs):
        arr[i].type = g.geometry.shape_type
        arr[i].parent_joint = g.parentJoint - 1
        arr[i].params[:] = g.geometry.params()
        arr[i].placement[:] = g.placement.rt34().reshape(-1).tolist()
        pts = getattr(g.geometry, "points", None)
        if pts is not None:
            arr[i].vert_offset, arr[i].vert_count = nv, len(pts)
            verts.append(pts)
            nv += len(pts)
    v = _f64(np.concatenate(verts, axis=0)) if verts else np.zeros((0, 3))
    return arr, v


class ModelDesc:
    """Owns the numpy/ctypes storage behind a ``RampModelDesc``."""

    def __init__(self, model: Model, geom_model: GeometryModel, n_geoms: int | None = None,
                 pairs=None):
        objs = geom_model.geometryObjects[: n_geoms if n_geoms is not None else None]
        pair_list = geom_model.collisionPairs if pairs is None else pairs
        nj = model.njoints - 1
        self.joints = (RampJoint * max(nj, 1))()
        for i in range(nj):
            j = model.joints[i + 1]
            self.joints[i].type = JOINT_TYPE[j.shortname()]
            self.joints[i].parent = model.parents[i + 1] - 1
            self.joints[i].idx_q = j.idx_q
            self.joints[i].axis[:] = np.asarray(j.axis, dtype=np.float64).tolist()
            self.joints[i].placement[:] = model.jointPlacements[i + 1].rt34().reshape(-1).tolist()
        self.geoms, self.verts = fill_geoms(objs)
        self.pairs = _i32([[p.first, p.second] for p in pair_list]).reshape(-1, 2)
        self.n_geoms = len(objs)
        self.struct = RampModelDesc(
            ABI_VERSION, model.nq, nj, C.cast(self.joints, C.POINTER(RampJoint)),
            len(objs), C.cast(self.geoms, C.POINTER(RampGeom)),
            len(self.pairs), _ptr(self.pairs, C.c_int32),
            len(self.verts), _ptr(self.verts, C.c_double))

    @property
    def n_pairs(self) -> int:
        return len(self.pairs)


class QMap:
    """Actuated columns -> Pinocchio q (mimic, then continuous); see RampQMap."""

    def __init__(self, nq: int, act_slot, ref_q, mimic_dst=(), mimic_src=(), mimic_mult=(),
                 mimic_off=(), cont_slot=()):
        self.act_slot = _i32(act_slot)
        self.mimic_dst, self.mimic_src = _i32(mimic_dst), _i32(mimic_src)
        self.mimic_mult, self.mimic_off = _f64(mimic_mult), _f64(mimic_off)
        self.cont_slot = _i32(cont_slot)
        self.ref_q = _f64(ref_q)
        assert len(self.ref_q) == nq
        self.nq = nq
        self.struct = RampQMap(
            len(self.act_slot), _ptr(self.act_slot, C.c_int32),
            len(self.mimic_dst), _ptr(self.mimic_dst, C.c_int32), _ptr(self.mimic_src, C.c_int32),
            _ptr(self.mimic_mult, C.c_double), _ptr(self.mimic_off, C.c_double),
            len(self.cont_slot), _ptr(self.cont_slot, C.c_int32), _ptr(self.ref_q, C.c_double))

    @property
    def n_act(self) -> int:
        return len(self.act_slot)

    def apply(self, act: np.ndarray) -> np.ndarray:
        """Host (numpy, float64) statement of the map -- used by the single-state API."""
        act = np.atleast_2d(np.asarray(act, dtype=np.float64))
        q = np.tile(self.ref_q, (len(act), 1))
        q[:, self.act_slot] = act
        if len(self.mimic_dst):
            q[:, self.mimic_dst] = q[:, self.mimic_src] * self.mimic_mult + self.mimic_off
        if len(self.cont_slot):
            th = q[:, self.cont_slot].copy()
            q[:, self.cont_slot] = np.cos(th)
            q[:, self.cont_slot + 1] = np.sin(th)
        return q


class IkParams:
    """Parameters of ramp_ik_batch for one tool frame (RampIkParams); keeps the bound arrays alive."""

    def __init__(self, frame_parent_joint: int, frame_placement, lower, upper, *, max_iterations: int = 500,
                 tolerance: float = 1e-3, damping: float = 0.05, step: float = 1.0, max_step: float = 0.5):
        self.lower, self.upper = _f64(lower), _f64(upper)
        pl = (C.c_double * 12)(*np.asarray(frame_placement, dtype=np.float64).reshape(12))
        self.struct = RampIkParams(int(frame_parent_joint), int(max_iterations), pl, float(damping), float(step),
                                   float(tolerance), float(max_step), _ptr(self.lower, C.c_double), _ptr(self.upper, C.c_double))


class SpaceDesc:
    """OMPL compound space over the actuated columns (reference motion_planner.py:264-359)."""

    def __init__(self, subspaces: list[dict], fraction: float = 0.01, check_bounds: bool = True):
        self.subspaces = subspaces
        self.arr = (RampSubspace * max(len(subspaces), 1))()
        for i, s in enumerate(subspaces):
            self.arr[i].type, self.arr[i].col, self.arr[i].dim = s["type"], s["col"], s.get("dim", 0)
            lo, hi = list(s.get("lo", [])), list(s.get("hi", []))
            self.arr[i].lo[:] = (lo + [0.0] * 3)[:3]
            self.arr[i].hi[:] = (hi + [0.0] * 3)[:3]
        self.struct = RampSpaceDesc(len(subspaces), C.cast(self.arr, C.POINTER(RampSubspace)),
                                    float(fraction), int(check_bounds), 0)

    @property
    def n_cols(self) -> int:
        width = {SPACE_SO2: 1, SPACE_SE2: 3, SPACE_SE3: 7, SPACE_SO3: 4}
        return max((s["col"] + (s["dim"] if s["type"] == SPACE_REAL else width[s["type"]])
                    for s in self.subspaces), default=0)


def is_unbounded(shortname: str) -> bool:
    return re.match(r"JointModelRUB([XYZ])$", shortname) is not None
