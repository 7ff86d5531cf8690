"""Host-side kinematic tree and geometry containers (setup-time only, never on the hot path).

They carry exactly the Pinocchio ``Model`` / ``GeometryModel`` facts the reference reads
(reference src/ramp/robot_model.py:53-55,130-139,328-339; src/ramp/pinocchio_utils.py:42-96;
src/ramp/compute_disable_collisions.py:38-53,200-226) with the same attribute names, so the
mirrored RobotModel / RobotState code and its tests read like the reference's.  The layout
conventions restated from Pinocchio 3.7.0: joint 0 is ``universe``,
``parents[i] < i``, fixed URDF joints create no joint but a FIXED_JOINT frame, a geometry's
placement is relative to its parent joint.
"""

from __future__ import annotations

import dataclasses
import re

import numpy as np

from ramp_b200.se3 import SE3
from ramp_b200.shapes import CollisionGeometry

SIZE_T_MAX = int(np.iinfo(np.uintp).max)

# Pinocchio FrameType values
FIXED_JOINT, JOINT, BODY, OP_FRAME, SENSOR = 1, 2, 4, 8, 16

# shortname -> (nq, nv)
JOINT_DIMS = {
    "JointModelRX": (1, 1), "JointModelRY": (1, 1), "JointModelRZ": (1, 1),
    "JointModelRevoluteUnaligned": (1, 1),
    "JointModelRUBX": (2, 1), "JointModelRUBY": (2, 1), "JointModelRUBZ": (2, 1),
    "JointModelRevoluteUnboundedUnaligned": (2, 1),
    "JointModelPX": (1, 1), "JointModelPY": (1, 1), "JointModelPZ": (1, 1),
    "JointModelPrismaticUnaligned": (1, 1),
    "JointModelPlanar": (4, 3),
    "JointModelFreeFlyer": (7, 6),
    "JointModelSpherical": (4, 3),
    "JointModelSphericalZYX": (3, 3),
    "JointModelTranslation": (3, 3),
}  # fmt: skip


@dataclasses.dataclass
class JointModel:
    """One movable joint (``model.joints[id]``)."""

    id: int
    _shortname: str
    idx_q: int
    idx_v: int
    axis: np.ndarray = dataclasses.field(default_factory=lambda: np.zeros(3))

    def shortname(self) -> str:
        return self._shortname

    @property
    def nq(self) -> int:
        return JOINT_DIMS[self._shortname][0] if self.id > 0 else 0

    @property
    def nv(self) -> int:
        return JOINT_DIMS[self._shortname][1] if self.id > 0 else 0


@dataclasses.dataclass
class Frame:
    name: str
    parentJoint: int  # noqa: N815
    parentFrame: int  # noqa: N815
    placement: SE3
    type: int


class Model:
    """Kinematic tree in Pinocchio order."""

    def __init__(self, name: str = "model"):
        self.name = name
        self.names: list[str] = ["universe"]
        self.joints: list[JointModel] = [JointModel(0, "JointModelUniverse", 0, 0)]
        self.parents: list[int] = [0]
        self.jointPlacements: list[SE3] = [SE3.Identity()]  # noqa: N815
        self.frames: list[Frame] = [Frame("universe", 0, 0, SE3.Identity(), FIXED_JOINT)]
        self.nq = 0
        self.nv = 0
        self._lower: list[float] = []
        self._upper: list[float] = []

    # -- construction ------------------------------------------------------
    def addJoint(self, parent: int, shortname: str, placement: SE3, name: str,  # noqa: N802
                 axis=None, lower=None, upper=None) -> int:
        if shortname not in JOINT_DIMS:
            raise ValueError(f"Unknown joint type: '{shortname}' for joint '{name}'")
        nq, nv = JOINT_DIMS[shortname]
        jid = len(self.joints)
        ax = np.zeros(3) if axis is None else np.asarray(axis, dtype=np.float64)
        self.joints.append(JointModel(jid, shortname, self.nq, self.nv, ax))
        self.names.append(name)
        self.parents.append(int(parent))
        self.jointPlacements.append(placement)
        lo, hi = self._default_limits(shortname)
        if lower is not None:
            lo = list(np.atleast_1d(lower).astype(float))
        if upper is not None:
            hi = list(np.atleast_1d(upper).astype(float))
        assert len(lo) == nq and len(hi) == nq
        self._lower.extend(lo)
        self._upper.extend(hi)
        self.nq += nq
        self.nv += nv
        return jid

    @staticmethod
    def _default_limits(shortname: str):
        inf = float("inf")
        cs = 1.01  # Pinocchio's bound on cos/sin and quaternion slots
        if re.match(r"JointModelRUB|JointModelRevoluteUnboundedUnaligned", shortname):
            return [-cs, -cs], [cs, cs]
        if shortname == "JointModelPlanar":
            return [-inf, -inf, -cs, -cs], [inf, inf, cs, cs]
        if shortname == "JointModelFreeFlyer":
            return [-inf] * 3 + [-cs] * 4, [inf] * 3 + [cs] * 4
        if shortname == "JointModelSpherical":
            return [-cs] * 4, [cs] * 4
        nq = JOINT_DIMS[shortname][0]
        return [-inf] * nq, [inf] * nq

    def addFrame(self, frame: Frame) -> int:  # noqa: N802
        self.frames.append(frame)
        return len(self.frames) - 1

    # -- queries (Pinocchio names) ----------------------------------------
    @property
    def njoints(self) -> int:
        return len(self.joints)

    @property
    def lowerPositionLimit(self) -> np.ndarray:  # noqa: N802
        return np.asarray(self._lower, dtype=np.float64)

    @property
    def upperPositionLimit(self) -> np.ndarray:  # noqa: N802
        return np.asarray(self._upper, dtype=np.float64)

    def getJointId(self, name: str) -> int:  # noqa: N802
        try:
            return self.names.index(name)
        except ValueError:
            return len(self.names)

    def existJointName(self, name: str) -> bool:  # noqa: N802
        return name in self.names

    def getFrameId(self, name: str, type: int | None = None) -> int:  # noqa: N802, A002
        for i, f in enumerate(self.frames):
            if f.name == name and (type is None or f.type & type):
                return i
        return len(self.frames)

    def getBodyId(self, name: str) -> int:  # noqa: N802
        return self.getFrameId(name, BODY)

    def existFrame(self, name: str) -> bool:  # noqa: N802
        return self.getFrameId(name) < len(self.frames)


def neutral(model: Model) -> np.ndarray:
    """pinocchio.neutral: zeros, (cos, sin) = (1, 0), quaternions = (0, 0, 0, 1)."""
    q = np.zeros(model.nq)
    for j in model.joints[1:]:
        s = j.shortname()
        if re.match(r"JointModelRUB|JointModelRevoluteUnboundedUnaligned", s):
            q[j.idx_q] = 1.0
        elif s == "JointModelPlanar":
            q[j.idx_q + 2] = 1.0
        elif s == "JointModelFreeFlyer":
            q[j.idx_q + 6] = 1.0
        elif s == "JointModelSpherical":
            q[j.idx_q + 3] = 1.0
    return q


def random_configuration(model: Model, rng: np.random.Generator | None = None,
                         position_bound: float = 1.0) -> np.ndarray:
    """pinocchio.randomConfiguration: uniform in the limits per bounded dof, uniform on the
    circle / unit-quaternion sphere for unbounded parts.  Unbounded translations (where
    Pinocchio would raise) are drawn from U[-position_bound, position_bound]."""
    rng = rng or np.random.default_rng()
    q = np.zeros(model.nq)
    lo, hi = model.lowerPositionLimit, model.upperPositionLimit

    def uni(i):
        a = lo[i] if np.isfinite(lo[i]) else -position_bound
        b = hi[i] if np.isfinite(hi[i]) else position_bound
        return rng.uniform(a, b)

    for j in model.joints[1:]:
        s, i = j.shortname(), j.idx_q
        if re.match(r"JointModelRUB|JointModelRevoluteUnboundedUnaligned", s):
            th = rng.uniform(-np.pi, np.pi)
            q[i], q[i + 1] = np.cos(th), np.sin(th)
        elif s == "JointModelPlanar":
            th = rng.uniform(-np.pi, np.pi)
            q[i], q[i + 1], q[i + 2], q[i + 3] = uni(i), uni(i + 1), np.cos(th), np.sin(th)
        elif s in ("JointModelFreeFlyer", "JointModelSpherical"):
            o = 3 if s == "JointModelFreeFlyer" else 0
            for k in range(o):
                q[i + k] = uni(i + k)
            v = rng.normal(size=4)
            q[i + o : i + o + 4] = v / np.linalg.norm(v)
        else:
            for k in range(j.nq):
                q[i + k] = uni(i + k)
    return q


class CollisionPair:
    __slots__ = ("first", "second")

    def __init__(self, first: int, second: int):
        self.first, self.second = int(first), int(second)

    def __eq__(self, other):
        return {self.first, self.second} == {other.first, other.second}

    def __hash__(self):
        return hash(frozenset((self.first, self.second)))

    def __repr__(self):
        return f"CollisionPair({self.first}, {self.second})"


class GeometryObject:
    """pinocchio.GeometryObject; accepts the 4- and 5-argument constructors RAMP uses
    (reference src/ramp/robot_model.py:453, examples/airplane_motion_planning.py:42-50)."""

    def __init__(self, name: str, parent_joint: int, *args):
        if len(args) == 2:
            parent_frame = SIZE_T_MAX
            a, b = args
        elif len(args) == 3:
            parent_frame, a, b = args
        else:
            raise TypeError("GeometryObject(name, parent_joint, [parent_frame,] placement, geometry)")
        # Pinocchio 3 takes (placement, geometry); Pinocchio 2 took (geometry, placement)
        placement, geometry = (a, b) if isinstance(a, SE3) else (b, a)
        if not isinstance(geometry, CollisionGeometry):
            raise TypeError("geometry must be a ramp_b200.shapes.CollisionGeometry")
        self.name = name
        self.parentJoint = int(parent_joint)  # noqa: N815
        self.parentFrame = int(parent_frame)  # noqa: N815
        self.placement = placement
        self.geometry = geometry
        self.meshColor = np.array([0.5, 0.5, 0.5, 1.0])  # noqa: N815
        self.disableCollision = False  # noqa: N815


class GeometryModel:
    def __init__(self):
        self.geometryObjects: list[GeometryObject] = []  # noqa: N815
        self.collisionPairs: list[CollisionPair] = []  # noqa: N815

    @property
    def ngeoms(self) -> int:
        return len(self.geometryObjects)

    def addGeometryObject(self, obj: GeometryObject) -> int:  # noqa: N802
        self.geometryObjects.append(obj)
        return len(self.geometryObjects) - 1

    def getGeometryId(self, name: str) -> int:  # noqa: N802
        for i, g in enumerate(self.geometryObjects):
            if g.name == name:
                return i
        return len(self.geometryObjects)

    def existGeometryName(self, name: str) -> bool:  # noqa: N802
        return self.getGeometryId(name) < len(self.geometryObjects)

    def addCollisionPair(self, pair: CollisionPair) -> None:  # noqa: N802
        n = len(self.geometryObjects)
        if pair.first == pair.second or not (0 <= pair.first < n and 0 <= pair.second < n):
            raise ValueError(f"invalid collision pair {pair}")
        if pair not in self.collisionPairs:
            self.collisionPairs.append(pair)

    def addAllCollisionPairs(self) -> None:  # noqa: N802
        """All i<j pairs whose parent joints differ, in (i, j) lexicographic order
        (pinocchio GeometryModel::addAllCollisionPairs; reference robot_model.py:328)."""
        self.collisionPairs = []
        n = len(self.geometryObjects)
        for i in range(n):
            for j in range(i + 1, n):
                if self.geometryObjects[i].parentJoint != self.geometryObjects[j].parentJoint:
                    self.collisionPairs.append(CollisionPair(i, j))

    def removeCollisionPair(self, pair: CollisionPair) -> None:  # noqa: N802
        if pair in self.collisionPairs:
            self.collisionPairs.remove(pair)

    def findCollisionPair(self, pair: CollisionPair) -> int:  # noqa: N802
        for i, p in enumerate(self.collisionPairs):
            if p == pair:
                return i
        return len(self.collisionPairs)

    def existCollisionPair(self, pair: CollisionPair) -> bool:  # noqa: N802
        return pair in self.collisionPairs


def remove_collision_pairs_from_srdf(model: Model, geom_model: GeometryModel,
                                     disabled: list[tuple[str, str]]) -> None:
    """pinocchio.removeCollisionPairsFromXML restated: for every
    ``<disable_collisions link1 link2>`` drop the pairs whose two geometries hang off those two
    BODY frames (keyed on frame ids, not geometry names; reference robot_model.py:330-339)."""
    for l1, l2 in disabled:
        f1, f2 = model.getBodyId(l1), model.getBodyId(l2)
        if f1 >= len(model.frames) or f2 >= len(model.frames):
            continue  # Pinocchio logs and skips unknown links
        keep = []
        for p in geom_model.collisionPairs:
            a = geom_model.geometryObjects[p.first].parentFrame
            b = geom_model.geometryObjects[p.second].parentFrame
            if (a == f1 and b == f2) or (a == f2 and b == f1):
                continue
            keep.append(p)
        geom_model.collisionPairs = keep
