"""Self-contained robot descriptions: RAMP's in-tree robots and stand-ins for the fetched ones.

Two families:

* ``rrr`` (+ planar / floating mobile base), ``planar_rrr``, ``rrr_mj``, ``acrobot``,
  ``airplane`` (floating / planar box) and ``double_pendulum`` restate, as generated
  URDF/MJCF/SRDF text, the facts of the reference's in-tree descriptions (reference
  robots/rrr/rrr.urdf.xacro, robots/rrr/rrr.srdf.xacro, robots/planar_rrr/robot.urdf.xacro,
  robots/rrr_mj/rrr.xml, robots/acrobot/acrobot.xml, robots/floating.urdf, robots/planar.urdf,
  tests/double_pendulum.urdf) and of their TOML configs (robots/*/configs.toml).  The text
  goes through ramp_b200.description, so the readers are exercised on every load;
  tests/test_description.py additionally loads the reference's own files when they are
  present and checks both routes give identical tables.

* ``panda``, ``ur5e``, ``fr3``, ``unitree_h1`` are STAND-IN MODELS.  The reference fetches
  those descriptions (URDF/MJCF + meshes) from the network through ``robot_descriptions``
  (reference src/ramp/pinocchio_utils.py:193-205); nothing of them exists offline.  Kinematic
  tables are the ones recorded in SURVEY App. A (from memory of the upstream assets), collision
  geometry is a procedurally generated 64-vertex convex hull per link, and the group / named
  state / disabled-pair data follow reference robots/{panda,ur5e,fr3,unitree_h1}/configs.toml
  and *.srdf.xacro.  Every result produced from them must be labelled "stand-in model".
"""

from __future__ import annotations

import numpy as np

from ramp_b200 import shapes
from ramp_b200.kinematics import BODY, FIXED_JOINT, JOINT, Frame, GeometryModel, GeometryObject, Model
from ramp_b200.se3 import SE3, rpy_to_matrix

PI = float(np.pi)


# ----------------------------------------------------------------- text emitters
def _link(name, geom=None, origin=None):
    if geom is None:
        return f'<link name="{name}"/>'
    o = f'<origin xyz="{origin}" rpy="0 0 0"/>' if origin else ""
    return f'<link name="{name}"><collision>{o}<geometry>{geom}</geometry></collision></link>'


def _joint(name, jtype, parent, child, xyz="0 0 0", axis=None, limit=None, mimic=None):
    s = f'<joint name="{name}" type="{jtype}"><parent link="{parent}"/><child link="{child}"/>'
    s += f'<origin xyz="{xyz}" rpy="0 0 0"/>'
    if axis:
        s += f'<axis xyz="{axis}"/>'
    if limit:
        s += f'<limit lower="{limit[0]}" upper="{limit[1]}" effort="100" velocity="1"/>'
    if mimic:
        s += f'<mimic joint="{mimic[0]}" multiplier="{mimic[1]}"/>'
    return s + "</joint>"


def _box(x, y, z):
    return f'<box size="{x} {y} {z}"/>'


def rrr_urdf(mobile_base_type: str = "") -> str:
    """Three revolute joints (z, y, y), box links, cylinder base, sphere tip; optional
    ``planar`` / ``floating`` mobile base under a ``world`` root."""
    lim = (-3.14159, 3.14159)
    parts = ['<robot name="rrr_robot">']
    if mobile_base_type:
        parts += [_link("world"), _link("mobile_base", _box(0.4, 0.3, 0.05))]
    parts += [
        _link("base_link", '<cylinder length="0.05" radius="0.1"/>'),
        _link("link1", _box(0.1, 0.1, 0.5), "0 0 0.25"),
        _link("link2", _box(0.1, 0.1, 0.4), "0 0 0.2"),
        _link("link3", _box(0.1, 0.1, 0.3), "0 0 0.15"),
        _link("end_effector", '<sphere radius="0.05"/>'),
    ]
    if mobile_base_type:
        parts += [
            _joint(f"{mobile_base_type}_joint", mobile_base_type, "world", "mobile_base"),
            _joint("base_mount", "fixed", "mobile_base", "base_link", "0 0 0.05"),
        ]
    parts += [
        _joint("joint1", "revolute", "base_link", "link1", "0 0 0.025", "0 0 1", lim),
        _joint("joint2", "revolute", "link1", "link2", "0 0 0.5", "0 1 0", lim),
        _joint("joint3", "revolute", "link2", "link3", "0 0 0.4", "0 1 0", lim),
        _joint("end_effector_joint", "fixed", "link3", "end_effector", "0 0 0.3", "0 0 1"),
        "</robot>",
    ]
    return "".join(parts)


def rrr_disabled(mobile_base_type: str = "") -> list[tuple[str, str]]:
    d = [("base_link", "link1"), ("link1", "link2"), ("link2", "link3"), ("link3", "end_effector")]
    return ([("mobile_base", "base_link")] if mobile_base_type else []) + d


def planar_rrr_urdf() -> str:
    parts = ['<robot name="planar_rrr_robot">', _link("base_link", '<cylinder length="0.1" radius="0.1"/>')]
    prev = "base_link"
    for k in (1, 2, 3):
        parts.append(_link(f"link{k}", _box(0.1, 0.1, 1.0), "0 0 0.5"))
        parts.append(_joint(f"joint{k}", "continuous", prev, f"link{k}",
                            "0 0 0" if k == 1 else "0 0 1.0", "1 0 0"))
        prev = f"link{k}"
    parts += [_link("end_effector"), _joint("end_effector_joint", "fixed", "link3", "end_effector", "0 0 1.0"),
              "</robot>"]
    return "".join(parts)


def airplane_urdf(kind: str) -> str:
    return "".join(['<robot name="airplane">', _link("world"), _link("mobile_base", _box(0.1, 0.075, 0.05)),
                    _joint(f"{kind}_joint", kind, "world", "mobile_base"), "</robot>"])


def double_pendulum_urdf() -> str:
    return "".join([
        '<robot name="2dof_planar">', _link("base_link"), _link("link1"), _link("link2"), _link("link3"),
        _joint("joint1", "continuous", "base_link", "link1", "0 0 0.035", "1 0 0"),
        _joint("joint2", "continuous", "link1", "link2", "0.023 0 0.1", "1 0 0"),
        _joint("joint3", "revolute", "link2", "link3", "0 0 0.1", "0 -1 0", (-3.14, 3.14), ("joint2", -0.5)),
        "</robot>"])


def acrobot_mjcf() -> str:
    return ('<mujoco model="simple_acrobot"><worldbody><body name="upper_link" pos="0 0 2.5">'
            '<joint axis="0 1 0" pos="0 0 0.75" name="shoulder" type="hinge"/>'
            '<geom name="upper_link" size="0.05 0.75" type="cylinder"/>'
            '<body name="lower_link" pos="0 0 -1.5">'
            '<joint axis="0 1 0" name="elbow" pos="0 0 0.75" type="hinge"/>'
            '<geom name="lower_link" size="0.05 0.75" type="cylinder"/>'
            "</body></body></worldbody></mujoco>")


def rrr_mjcf() -> str:
    return ('<mujoco model="rrr_robot"><default><default class="joint"><joint range="-3.14 3.14"/></default></default>'
            '<worldbody><body name="base_link" pos="0 0 0"><geom type="cylinder" size="0.1 0.025"/>'
            '<body name="link1" pos="0 0 0.025"><joint name="joint1" class="joint" axis="0 0 1"/>'
            '<geom type="box" size="0.05 0.05 0.25" pos="0 0 0.25"/>'
            '<body name="link2" pos="0 0 0.5"><joint name="joint2" class="joint" axis="0 1 0"/>'
            '<geom type="box" size="0.05 0.05 0.2" pos="0 0 0.2"/>'
            '<body name="link3" pos="0 0 0.4"><joint name="joint3" class="joint" axis="0 1 0"/>'
            '<geom type="box" size="0.05 0.05 0.15" pos="0 0 0.15"/>'
            '<body name="end_effector" pos="0 0 0.3"><geom type="sphere" size="0.05"/></body>'
            "</body></body></body></body></worldbody></mujoco>")


# ----------------------------------------------------------- stand-in geometry
def capsule_hull(p0, p1, radius: float, seed: int, n_ring: int = 16) -> np.ndarray:
    """64 vertices: a ring of 16 at each end of the segment p0-p1 plus a 15-vertex cap ring
    and an apex beyond each end; a small seeded radial jitter keeps the hull generic."""
    rng = np.random.default_rng(seed)
    p0, p1 = np.asarray(p0, float), np.asarray(p1, float)
    ax = p1 - p0
    length = np.linalg.norm(ax)
    ez = ax / length if length > 1e-9 else np.array([0.0, 0.0, 1.0])
    tmp = np.array([1.0, 0.0, 0.0]) if abs(ez[0]) < 0.9 else np.array([0.0, 1.0, 0.0])
    ex = np.cross(ez, tmp)
    ex /= np.linalg.norm(ex)
    ey = np.cross(ez, ex)
    pts = []
    for centre, sign in ((p0, -1.0), (p1, 1.0)):
        for k in range(n_ring):
            a = 2 * PI * (k + 0.5 * (sign > 0)) / n_ring
            r = radius * (1.0 + 0.06 * rng.uniform(-1, 1))
            pts.append(centre + r * (np.cos(a) * ex + np.sin(a) * ey))
        for k in range(n_ring - 1):
            a = 2 * PI * k / (n_ring - 1)
            r = 0.62 * radius * (1.0 + 0.06 * rng.uniform(-1, 1))
            pts.append(centre + r * (np.cos(a) * ex + np.sin(a) * ey) + sign * 0.72 * radius * ez)
        pts.append(centre + sign * radius * ez)
    return np.asarray(pts)


def _segment_frame(p0, p1):
    p0, p1 = np.asarray(p0, float), np.asarray(p1, float)
    ax = p1 - p0
    length = float(np.linalg.norm(ax))
    ez = ax / length if length > 1e-9 else np.array([0.0, 0.0, 1.0])
    ex = np.cross(ez, np.array([1.0, 0.0, 0.0]) if abs(ez[0]) < 0.9 else np.array([0.0, 1.0, 0.0]))
    ex /= np.linalg.norm(ex)
    return 0.5 * (p0 + p1), np.column_stack([ex, np.cross(ez, ex), ez]), length


def _superellipsoid(half, n: int, power: float, rng) -> np.ndarray:
    """n points on |x/a|^p + |y/b|^p + |z/c|^p = 1 along quasi-uniform directions (p = 2: an
    ellipsoid, p >= 6: a box with rounded edges).  The body is convex, so every point is a
    vertex of the hull of the set."""
    k = np.arange(n) + 0.5
    phi = np.arccos(1.0 - 2.0 * k / n)
    th = np.pi * (1.0 + 5.0 ** 0.5) * k + rng.uniform(0, 2 * np.pi)
    d = np.column_stack([np.cos(th) * np.sin(phi), np.sin(th) * np.sin(phi), np.cos(phi)])
    d += 0.02 * rng.normal(size=d.shape)
    half = np.asarray(half, float)
    scale = (np.abs(d / half) ** power).sum(axis=1) ** (-1.0 / power)
    return d * scale[:, None]


def shaped_hull(kind: str, p0, p1, radius: float, seed: int) -> np.ndarray:
    """Stand-in link geometry that is NOT capsule-like, around the segment p0-p1 (the same
    segments and radii the capsule-like stand-ins use).  Real link meshes are boxy housings,
    elbows and flat plates; an inscribed capsule fills little of such a hull, so far more
    (configuration, pair) items fall between the inner and the outer capsule and reach GJK.

    kind: ``rbox``  a box with rounded edges, 64 vertices
          ``L``     the hull of two rounded boxes at a right angle (an elbow housing), <= 80
          ``plate`` a flat slab, 64 vertices
          ``blob``  an irregular ellipsoid-like body, ~100-128 vertices
    """
    from ramp_b200 import meshes

    rng = np.random.default_rng(seed)
    centre, R, length = _segment_frame(p0, p1)
    radius = 0.9 * radius  # corners reach further than a capsule of the same radius would
    hz = 0.5 * length + radius
    if kind == "rbox":
        local = _superellipsoid((radius, 0.8 * radius, hz), 64, 6.0, rng)
    elif kind == "plate":
        local = _superellipsoid((1.3 * radius, 0.4 * radius, hz), 64, 8.0, rng)
    elif kind == "L":
        a = _superellipsoid((0.85 * radius, 0.85 * radius, hz), 40, 6.0, rng)
        b = _superellipsoid((1.5 * radius, 0.8 * radius, 0.8 * radius), 40, 6.0, rng)
        b += np.array([1.1 * radius, 0.0, hz - 0.8 * radius])
        local = meshes.convex_hull_vertices(np.concatenate([a, b]), 128)
    elif kind == "blob":
        pts = _superellipsoid((1.05 * radius, 0.9 * radius, hz), 128, 2.0, rng)
        pts *= (1.0 + 0.05 * rng.uniform(-1, 1, size=(len(pts), 1)))
        local = meshes.convex_hull_vertices(pts, 128)
    else:
        raise ValueError(f"unknown stand-in hull kind '{kind}'")
    return centre + local @ R.T


class _Builder:
    """Programmatic Model/GeometryModel construction with Pinocchio's frame bookkeeping."""

    def __init__(self, name: str, root_link: str):
        self.model = Model(name)
        self.geom = GeometryModel()
        self.body = {root_link: self.model.addFrame(Frame(root_link, 0, 0, SE3.Identity(), BODY))}
        self.mimics: list[dict] = []

    def joint(self, name, parent_link, child_link, short, xyz, rpy=(0, 0, 0), axis=None, lower=None, upper=None):
        m = self.model
        pf = m.frames[self.body[parent_link]]
        placement = pf.placement * SE3(rpy_to_matrix(*rpy), xyz)
        jid = m.addJoint(pf.parentJoint, short, placement, name, axis=axis, lower=lower, upper=upper)
        fid = m.addFrame(Frame(name, jid, self.body[parent_link], SE3.Identity(), JOINT))
        self.body[child_link] = m.addFrame(Frame(child_link, jid, fid, SE3.Identity(), BODY))

    def fixed(self, name, parent_link, child_link, xyz, rpy=(0, 0, 0)):
        m = self.model
        pf = m.frames[self.body[parent_link]]
        placement = pf.placement * SE3(rpy_to_matrix(*rpy), xyz)
        fid = m.addFrame(Frame(name, pf.parentJoint, self.body[parent_link], placement, FIXED_JOINT))
        self.body[child_link] = m.addFrame(Frame(child_link, pf.parentJoint, fid, placement, BODY))

    def hull(self, link, p0, p1, radius, seed):
        """Attach a convex hull, re-centred so the local origin is its bounding-sphere
        centre (the placement absorbs the shift).  ``self.kinds`` (link -> kind of
        ``shaped_hull``) replaces the capsule-like hull for the "boxy" variants."""
        kind = getattr(self, "kinds", {}).get(link)
        pts = capsule_hull(p0, p1, radius, seed) if kind is None else shaped_hull(kind, p0, p1, radius, seed)
        centre = 0.5 * (pts.min(axis=0) + pts.max(axis=0))
        f = self.model.frames[self.body[link]]
        k = sum(1 for g in self.geom.geometryObjects if g.parentFrame == self.body[link])
        self.geom.addGeometryObject(GeometryObject(
            f"{link}_{k}", f.parentJoint, self.body[link], f.placement * SE3(np.eye(3), centre),
            shapes.Convex(pts - centre)))


    def prim(self, link, shape, xyz=(0.0, 0.0, 0.0), rpy=(0.0, 0.0, 0.0)):
        """Attach a primitive shape at an offset in the link frame."""
        f = self.model.frames[self.body[link]]
        k = sum(1 for g in self.geom.geometryObjects if g.parentFrame == self.body[link])
        self.geom.addGeometryObject(GeometryObject(
            f"{link}_{k}", f.parentJoint, self.body[link], f.placement * SE3(rpy_to_matrix(*rpy), xyz), shape))


def _joint_zoo():
    """One joint of EVERY type of the vocabulary (include/ramp_b200.h RampJointType; reference
    src/ramp/constants.py:14-23, src/ramp/motion_planner.py:268-362) in one tree with two
    branches, placements with rotations, a primitive on most links.  Not a robot: a test
    article for the forward-kinematics code of the rarely used joint types."""
    b = _Builder("joint_zoo", "root")
    ua = np.array([1.0, 2.0, -0.5])
    ua /= np.linalg.norm(ua)
    ub = np.array([-0.3, 0.4, 0.85])
    ub /= np.linalg.norm(ub)
    chain = [  # name, type, xyz, rpy, axis, lower, upper
        ("j_free", "JointModelFreeFlyer", (0, 0, 0), (0, 0, 0), None, None, None),
        ("j_rx", "JointModelRX", (0.1, 0, 0.05), (0.2, 0, 0), None, -2.0, 2.0),
        ("j_ry", "JointModelRY", (0, 0.1, 0.1), (0, 0.3, 0), None, -2.0, 2.0),
        ("j_rz", "JointModelRZ", (0, 0, 0.15), (0, 0, -0.4), None, -2.0, 2.0),
        ("j_ru", "JointModelRevoluteUnaligned", (0.05, 0, 0.1), (0.1, 0.2, 0.3), ua, -2.5, 2.5),
        ("j_rubx", "JointModelRUBX", (0, 0.05, 0.1), (0, 0, 0), None, None, None),
        ("j_ruby", "JointModelRUBY", (0, 0, 0.12), (0.3, 0, 0), None, None, None),
        ("j_rubz", "JointModelRUBZ", (0.04, 0, 0.1), (0, -0.2, 0), None, None, None),
        ("j_rubu", "JointModelRevoluteUnboundedUnaligned", (0, 0, 0.1), (0, 0.1, 0.5), ub, None, None),
    ]
    prev = "root"
    for k, (name, short, xyz, rpy, axis, lo, hi) in enumerate(chain):
        link = f"l_{name[2:]}"
        b.joint(name, prev, link, short, xyz, rpy, axis=axis, lower=lo, upper=hi)
        b.prim(link, shapes.Sphere(0.03) if k % 2 else shapes.Box(0.05, 0.04, 0.06), (0.0, 0.0, 0.03))
        prev = link
    # second branch off the free-flying body: prismatic, planar, spherical, ZYX, translation
    branch = [
        ("j_px", "JointModelPX", (-0.1, 0, 0), (0, 0, 0.3), None, -0.2, 0.2),
        ("j_py", "JointModelPY", (0, -0.1, 0.05), (0.2, 0, 0), None, -0.2, 0.2),
        ("j_pz", "JointModelPZ", (0, 0, -0.1), (0, 0.2, 0), None, -0.2, 0.2),
        ("j_pu", "JointModelPrismaticUnaligned", (-0.05, 0, -0.05), (0.1, 0.1, 0), ub, -0.2, 0.2),
        ("j_planar", "JointModelPlanar", (0, -0.05, -0.1), (0.4, 0, 0), None, None, None),
        ("j_sph", "JointModelSpherical", (0, 0, -0.12), (0, 0, 0.2), None, None, None),
        ("j_zyx", "JointModelSphericalZYX", (0.03, 0, -0.1), (0, 0.3, 0), None, [-1.5] * 3, [1.5] * 3),
        ("j_tr", "JointModelTranslation", (0, 0.03, -0.1), (0.2, 0, 0.1), None, [-0.15] * 3, [0.15] * 3),
    ]
    prev = "l_free"
    for k, (name, short, xyz, rpy, axis, lo, hi) in enumerate(branch):
        link = f"l_{name[2:]}"
        b.joint(name, prev, link, short, xyz, rpy, axis=axis, lower=lo, upper=hi)
        b.prim(link, shapes.Capsule(0.02, 0.06) if k % 2 else shapes.Cylinder(0.03, 0.05), (0.0, 0.0, -0.03), (0.3, 0.0, 0.0))
        prev = link
    return b


# which non-capsule shape each link of the "boxy" variants gets (shaped_hull kinds)
_PANDA_BOXY = {"link0": "rbox", "link1": "L", "link2": "blob", "link3": "L", "link4": "rbox", "link5": "L", "link6": "blob",
               "link7": "rbox", "hand": "plate", "leftfinger": "rbox", "rightfinger": "rbox"}
_H1_BOXY = {"pelvis": "rbox", "torso_link": "rbox", "hip_yaw_link": "blob", "hip_roll_link": "rbox", "hip_pitch_link": "rbox",
            "knee_link": "L", "ankle_link": "plate", "shoulder_pitch_link": "blob", "shoulder_roll_link": "rbox",
            "shoulder_yaw_link": "L", "elbow_link": "rbox"}


def _panda_like(prefix: str, with_hand: bool, name: str, boxy: bool = False):
    b = _Builder(name, f"{prefix}link0")
    if boxy:
        b.kinds = {prefix + k: v for k, v in _PANDA_BOXY.items()}
    h = PI / 2
    R = "JointModelRZ"
    J = [  # xyz, rpy, lower, upper  
        ((0, 0, 0.333), (0, 0, 0), -2.8973, 2.8973),
        ((0, 0, 0), (-h, 0, 0), -1.7628, 1.7628),
        ((0, -0.316, 0), (h, 0, 0), -2.8973, 2.8973),
        ((0.0825, 0, 0), (h, 0, 0), -3.0718, -0.0698),
        ((-0.0825, 0.384, 0), (-h, 0, 0), -2.8973, 2.8973),
        ((0, 0, 0), (h, 0, 0), -0.0175, 3.7525),
        ((0.088, 0, 0), (h, 0, 0), -2.8973, 2.8973),
    ]
    for k, (xyz, rpy, lo, hi) in enumerate(J, start=1):
        b.joint(f"{prefix}joint{k}", f"{prefix}link{k - 1}", f"{prefix}link{k}", R, xyz, rpy, lower=lo, upper=hi)
    H = [  # link -> segment end points in the link frame, radius
        ((-0.06, 0, 0.06), (0.0, 0, 0.08), 0.085),
        ((0, 0, -0.17), (0, -0.01, -0.02), 0.062),
        ((0, -0.02, 0), (0, -0.13, 0.01), 0.062),
        ((0, 0.005, -0.13), (0.07, 0.02, -0.02), 0.058),
        ((-0.015, 0.0, 0.0), (-0.075, 0.09, 0.01), 0.058),
        ((0, 0.04, -0.25), (0, 0.01, -0.04), 0.052),
        ((0.0, 0.0, 0.005), (0.075, 0.01, 0.005), 0.05),
        ((0, 0, 0.03), (0.01, 0.01, 0.075), 0.042),
    ]
    for k, (p0, p1, r) in enumerate(H):
        b.hull(f"{prefix}link{k}", p0, p1, r, seed=100 + k)
    if with_hand:
        b.fixed(f"{prefix}joint8", f"{prefix}link7", f"{prefix}link8", (0, 0, 0.107))
        b.fixed(f"{prefix}hand_joint", f"{prefix}link8", f"{prefix}hand", (0, 0, 0), (0, 0, -PI / 4))
        b.hull(f"{prefix}hand", (0, -0.07, 0.03), (0, 0.07, 0.03), 0.03, seed=120)
        b.joint(f"{prefix}finger_joint1", f"{prefix}hand", f"{prefix}leftfinger", "JointModelPY",
                (0, 0, 0.0584), lower=0.0, upper=0.04)
        b.joint(f"{prefix}finger_joint2", f"{prefix}hand", f"{prefix}rightfinger", "JointModelPrismaticUnaligned",
                (0, 0, 0.0584), axis=(0, -1, 0), lower=0.0, upper=0.04)
        b.mimics.append({"joint": f"{prefix}finger_joint2", "mimicked": f"{prefix}finger_joint1",
                         "multiplier": 1.0, "offset": 0.0})
        b.hull(f"{prefix}leftfinger", (0, 0.008, 0.012), (0, 0.008, 0.042), 0.011, seed=121)
        b.hull(f"{prefix}rightfinger", (0, -0.008, 0.012), (0, -0.008, 0.042), 0.011, seed=122)
        b.fixed(f"{prefix}hand_tcp_joint", f"{prefix}hand", f"{prefix}hand_tcp", (0, 0, 0.1034))
    else:
        b.fixed("attachment_joint", f"{prefix}link7", "attachment_site", (0, 0, 0.107))
    return b


# disabled pairs of reference robots/panda/panda.srdf.xacro:6-49 (prefix "panda_")
_PANDA_DISABLED = [
    ("link0", "link1"), ("link0", "link2"), ("link0", "link3"), ("link0", "link4"), ("link1", "link2"),
    ("link1", "link3"), ("link1", "link4"), ("link2", "link3"), ("link2", "link4"), ("link2", "link6"),
    ("link3", "link4"), ("link3", "link5"), ("link3", "link6"), ("link3", "link7"), ("link4", "link5"),
    ("link4", "link6"), ("link4", "link7"), ("link4", "link8"), ("link5", "link6"), ("link6", "link7"),
    ("link6", "link8"), ("link7", "link8"),
    ("hand", "leftfinger"), ("hand", "rightfinger"), ("leftfinger", "rightfinger"),
    ("hand", "link3"), ("hand", "link4"), ("hand", "link6"), ("hand", "link7"), ("hand", "link8"),
    ("leftfinger", "link3"), ("leftfinger", "link4"), ("leftfinger", "link6"), ("leftfinger", "link7"),
    ("leftfinger", "link8"), ("link3", "rightfinger"), ("link4", "rightfinger"), ("link6", "rightfinger"),
    ("link7", "rightfinger"), ("link8", "rightfinger"),
]  # fmt: skip


def _ur5():
    b = _Builder("ur5", "world")
    h = PI / 2
    tw = 2 * PI
    b.fixed("world_joint", "world", "base_link", (0, 0, 0))
    b.joint("shoulder_pan_joint", "base_link", "shoulder_link", "JointModelRZ", (0, 0, 0.089159), lower=-tw, upper=tw)
    b.joint("shoulder_lift_joint", "shoulder_link", "upper_arm_link", "JointModelRY", (0, 0.13585, 0), (0, h, 0), lower=-tw, upper=tw)
    b.joint("elbow_joint", "upper_arm_link", "forearm_link", "JointModelRY", (0, -0.1197, 0.425), lower=-tw, upper=tw)
    b.joint("wrist_1_joint", "forearm_link", "wrist_1_link", "JointModelRY", (0, 0, 0.39225), (0, h, 0), lower=-tw, upper=tw)
    b.joint("wrist_2_joint", "wrist_1_link", "wrist_2_link", "JointModelRZ", (0, 0.093, 0), lower=-tw, upper=tw)
    b.joint("wrist_3_joint", "wrist_2_link", "wrist_3_link", "JointModelRY", (0, 0, 0.09465), lower=-tw, upper=tw)
    b.fixed("ee_fixed_joint", "wrist_3_link", "ee_link", (0, 0.0823, 0), (0, 0, h))
    b.fixed("wrist_3_link-tool0_fixed_joint", "wrist_3_link", "tool0", (0, 0.0823, 0), (-h, 0, 0))
    H = {
        "base_link": ((0, 0, 0.01), (0, 0, 0.05), 0.075),
        "shoulder_link": ((0, 0, -0.02), (0, 0.06, 0.0), 0.065),
        "upper_arm_link": ((0, 0, 0.03), (0, 0, 0.40), 0.058),
        "forearm_link": ((0, 0, 0.03), (0, 0, 0.37), 0.045),
        "wrist_1_link": ((0, 0.03, 0), (0, 0.09, 0), 0.042),
        "wrist_2_link": ((0, 0, 0.03), (0, 0, 0.09), 0.042),
        "wrist_3_link": ((0, 0.02, 0), (0, 0.065, 0), 0.038),
        "ee_link": ((-0.01, 0, 0), (0.0, 0, 0), 0.012),
    }
    for k, (link, (p0, p1, r)) in enumerate(H.items()):
        b.hull(link, p0, p1, r, seed=200 + k)
    return b


# reference robots/ur5e/ur.srdf.xacro:3-10
_UR5_DISABLED = [("shoulder_link", "upper_arm_link"), ("wrist_2_link", "wrist_3_link"),
                 ("upper_arm_link", "forearm_link"), ("wrist_1_link", "wrist_2_link"),
                 ("base_link", "shoulder_link"), ("wrist_3_link", "ee_link"), ("world", "base_link"),
                 ("forearm_link", "wrist_1_link")]


def _h1(boxy: bool = False):
    """Floating base + 19 hinges; 20 geometry-bearing links (stand-in table, from memory of
    the upstream MJCF: left leg, right leg, torso, left arm, right arm in file order)."""
    b = _Builder("h1", "world_root")
    if boxy:
        b.kinds = {side + k: v for k, v in _H1_BOXY.items() for side in ("", "left_", "right_")}
    b.joint("floating_base_joint", "world_root", "pelvis", "JointModelFreeFlyer", (0, 0, 0))
    b.hull("pelvis", (0, -0.06, -0.03), (0, 0.06, -0.03), 0.09, seed=300)
    seed = 301
    for side, s in (("left", 1.0), ("right", -1.0)):
        b.joint(f"{side}_hip_yaw", "pelvis", f"{side}_hip_yaw_link", "JointModelRZ", (0, s * 0.0875, -0.1742), lower=-0.43, upper=0.43)
        b.joint(f"{side}_hip_roll", f"{side}_hip_yaw_link", f"{side}_hip_roll_link", "JointModelRX", (0.039468, 0, 0), lower=-0.43, upper=0.43)
        b.joint(f"{side}_hip_pitch", f"{side}_hip_roll_link", f"{side}_hip_pitch_link", "JointModelRY", (0, s * 0.11536, 0), lower=-1.57, upper=1.57)
        b.joint(f"{side}_knee", f"{side}_hip_pitch_link", f"{side}_knee_link", "JointModelRY", (0, 0, -0.4), lower=-0.26, upper=2.05)
        b.joint(f"{side}_ankle", f"{side}_knee_link", f"{side}_ankle_link", "JointModelRY", (0, 0, -0.4), lower=-0.87, upper=0.52)
        for link, p0, p1, r in (
            (f"{side}_hip_yaw_link", (-0.02, 0, 0), (0.02, 0, 0), 0.05),
            (f"{side}_hip_roll_link", (0.02, 0, 0), (0.02, s * 0.03, 0), 0.04),
            (f"{side}_hip_pitch_link", (0, 0, -0.10), (0, 0, -0.33), 0.06),
            (f"{side}_knee_link", (0, 0, -0.06), (0, 0, -0.34), 0.05),
            (f"{side}_ankle_link", (-0.05, 0, -0.045), (0.13, 0, -0.045), 0.03),
        ):
            b.hull(link, p0, p1, r, seed=seed)
            seed += 1
    b.joint("torso", "pelvis", "torso_link", "JointModelRZ", (0, 0, 0), lower=-2.35, upper=2.35)
    b.hull("torso_link", (0, 0, 0.12), (0, 0, 0.42), 0.12, seed=seed)
    seed += 1
    for side, s in (("left", 1.0), ("right", -1.0)):
        b.joint(f"{side}_shoulder_pitch", "torso_link", f"{side}_shoulder_pitch_link", "JointModelRY",
                (0.0055, s * 0.15535, 0.42999), (s * 0.43633, 0, 0), lower=-2.87, upper=2.87)
        b.joint(f"{side}_shoulder_roll", f"{side}_shoulder_pitch_link", f"{side}_shoulder_roll_link", "JointModelRX",
                (-0.0055, s * 0.0565, -0.0165), (-s * 0.43633, 0, 0),
                lower=-0.34 if s > 0 else -3.11, upper=3.11 if s > 0 else 0.34)
        b.joint(f"{side}_shoulder_yaw", f"{side}_shoulder_roll_link", f"{side}_shoulder_yaw_link", "JointModelRZ",
                (0, 0, -0.1343), lower=-1.3 if s > 0 else -4.45, upper=4.45 if s > 0 else 1.3)
        b.joint(f"{side}_elbow", f"{side}_shoulder_yaw_link", f"{side}_elbow_link", "JointModelRY",
                (0.0185, 0, -0.198), lower=-1.25, upper=2.61)
        for link, p0, p1, r in (
            (f"{side}_shoulder_pitch_link", (0, s * 0.01, 0), (0, s * 0.05, -0.01), 0.045),
            (f"{side}_shoulder_roll_link", (0, 0, -0.02), (0, 0, -0.09), 0.045),
            (f"{side}_shoulder_yaw_link", (0, 0, -0.04), (0.01, 0, -0.17), 0.042),
            (f"{side}_elbow_link", (0.03, 0, -0.01), (0.26, 0, -0.02), 0.04),
        ):
            b.hull(link, p0, p1, r, seed=seed)
            seed += 1
    return b


# reference robots/unitree_h1/h1.srdf.xacro:3-15
_H1_DISABLED = [
    ("pelvis", "torso_link"), ("left_hip_yaw_link", "left_hip_roll_link"),
    ("left_hip_pitch_link", "left_knee_link"), ("right_hip_yaw_link", "right_hip_roll_link"),
    ("right_hip_pitch_link", "right_knee_link"), ("torso_link", "left_shoulder_pitch_link"),
    ("torso_link", "right_shoulder_pitch_link"), ("left_shoulder_pitch_link", "left_shoulder_roll_link"),
    ("left_shoulder_roll_link", "left_shoulder_yaw_link"), ("left_shoulder_yaw_link", "left_elbow_link"),
    ("right_shoulder_pitch_link", "right_shoulder_roll_link"),
    ("right_shoulder_roll_link", "right_shoulder_yaw_link"), ("right_shoulder_yaw_link", "right_elbow_link"),
]

# reference robots/fr3/fr3.srdf.xacro:3-10
_FR3_DISABLED = [("fr3_link3", "fr3_link4"), ("fr3_link0", "fr3_link1"), ("fr3_link2", "fr3_link3"),
                 ("fr3_link5", "fr3_link6"), ("fr3_link6", "fr3_link7"), ("base", "fr3_link0"),
                 ("fr3_link4", "fr3_link5"), ("fr3_link1", "fr3_link2")]


def standin(description: str):
    """``robot-descriptions::<name>`` -> (Model, GeometryModel, mimics) stand-in, or None."""
    if description in ("panda_description", "panda_mj_description"):
        b = _panda_like("panda_", True, "panda")
    elif description in ("fr3_mj_description", "fr3_description"):
        b = _panda_like("fr3_", False, "fr3")
    elif description in ("ur5_description", "ur5e_description"):
        b = _ur5()
    elif description in ("h1_mj_description", "h1_description"):
        b = _h1()
    elif description == "joint_zoo":
        b = _joint_zoo()
    elif description == "panda_boxy_description":
        b = _panda_like("panda_", True, "panda", boxy=True)
    elif description == "h1_boxy_description":
        b = _h1(boxy=True)
    else:
        return None
    return b.model, b.geom, b.mimics


_ARM7 = [f"joint{k}" for k in range(1, 8)]
_ACC1 = {"joint1": 1.0, "joint2": 1.0, "joint3": 1.0}

# name -> configuration; the dict layout is the reference's TOML schema
# (robots/*/configs.toml) plus "text"/"kind"/"disabled" for the generated description.
CONFIGS: dict[str, dict] = {
    "rrr/configs": {
        "text": rrr_urdf(), "kind": "urdf", "disabled": rrr_disabled(),
        "robot": {"base_link": "base_link"},
        "group": {"arm": {"joints": ["joint1", "joint2", "joint3"], "tcp_link_name": "end_effector",
                          "named_states": {"home": [0.0, 0.0, 0.0]}}},
        "acceleration_limits": dict(_ACC1),
    },
    "rrr/planar_configs": {
        "text": rrr_urdf("planar"), "kind": "urdf", "disabled": rrr_disabled("planar"),
        "robot": {"base_link": "base_link"},
        "group": {"arm": {"joints": ["planar_joint", "joint1", "joint2", "joint3"], "tcp_link_name": "end_effector",
                          "named_states": {"home": [0.0] * 6}}},
        "acceleration_limits": {**_ACC1, "planar_joint": 1.0},
    },
    "rrr/floating_configs": {
        "text": rrr_urdf("floating"), "kind": "urdf", "disabled": rrr_disabled("floating"),
        "robot": {"base_link": "base_link"},
        "group": {"arm": {"joints": ["floating_joint", "joint1", "joint2", "joint3"], "tcp_link_name": "end_effector",
                          "named_states": {"home": [0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 1.0, 0.0, 0.0, 0.0]}}},
        "acceleration_limits": {**_ACC1, "floating_joint": 1.0},
    },
    "planar_rrr/configs": {
        "text": planar_rrr_urdf(), "kind": "urdf", "disabled": [],
        "robot": {"base_link": "base_link"},
        "group": {"arm": {"joints": ["joint1", "joint2", "joint3"], "tcp_link_name": "end_effector",
                          "named_states": {"home": [0.0, 0.0, 0.0]}}},
        "acceleration_limits": dict(_ACC1),
    },
    "rrr_mj/configs": {
        "text": rrr_mjcf(), "kind": "mjcf", "disabled": [],
        "robot": {"base_link": "base_link"},
        "group": {"arm": {"joints": ["joint1", "joint2", "joint3"], "tcp_link_name": "end_effector",
                          "named_states": {"home": [0.0, 0.0, 0.0]}}},
        "acceleration_limits": dict(_ACC1),
    },
    "acrobot/configs": {
        "text": acrobot_mjcf(), "kind": "mjcf", "disabled": [("upper_link", "lower_link")],
        "robot": {"base_link": "universe"},
        "group": {"arm": {"joints": ["elbow"], "named_states": {"home": [0.0]}}},
        "acceleration_limits": {"elbow": 1.0},
    },
    "double_pendulum": {
        "text": double_pendulum_urdf(), "kind": "urdf", "disabled": [],
        "robot": {"base_link": "base_link"},
        "group": {"arm": {"joints": ["joint1", "joint2"]}},
        "acceleration_limits": {"joint1": 1.0, "joint2": 1.0},
    },
    "floating": {"text": airplane_urdf("floating"), "kind": "urdf", "disabled": [], "bare": True},
    "planar": {"text": airplane_urdf("planar"), "kind": "urdf", "disabled": [], "bare": True},
    "panda/configs": {
        "standin": "panda_description", "disabled": [(f"panda_{a}", f"panda_{b}") for a, b in _PANDA_DISABLED],
        "robot": {"base_link": "panda_link0"},
        "group": {"arm": {"joints": [f"panda_{j}" for j in _ARM7], "tcp_link_name": "panda_hand_tcp",
                          "named_states": {"home": [0.0, -0.785, 0.0, -2.356, 0.0, 1.571, 0.785]}},
                  "gripper": {"joints": ["panda_finger_joint1"]}},
        "acceleration_limits": {**{f"panda_joint{k}": v for k, v in zip(range(1, 8), (3.75, 1.875, 2.5, 3.125, 3.75, 5.0, 5.0))},
                                "panda_finger_joint1": 1.0},
    },
    "fr3/configs": {
        "standin": "fr3_mj_description", "disabled": _FR3_DISABLED,
        "robot": {"base_link": "fr3_link0"},
        "group": {"arm": {"joints": [f"fr3_{j}" for j in _ARM7], "tcp_link_name": "attachment_site",
                          "named_states": {"home": [0.0, -0.785, 0.0, -2.356, 0.0, 1.571, 0.785]}}},
        "acceleration_limits": {f"fr3_joint{k}": v for k, v in zip(range(1, 8), (3.75, 1.875, 2.5, 3.125, 3.75, 5.0, 5.0))},
    },
    "ur5e/configs": {
        "standin": "ur5_description", "disabled": _UR5_DISABLED,
        "robot": {"base_link": "world"},
        "group": {"arm": {"joints": ["shoulder_pan_joint", "shoulder_lift_joint", "elbow_joint", "wrist_1_joint",
                                     "wrist_2_joint", "wrist_3_joint"], "tcp_link_name": "tool0",
                          "named_states": {"home": [0.0] * 6}}},
        "acceleration_limits": {j: 1.0 for j in ("shoulder_pan_joint", "shoulder_lift_joint", "elbow_joint",
                                                  "wrist_1_joint", "wrist_2_joint", "wrist_3_joint")},
    },
    "joint_zoo": {"standin": "joint_zoo", "disabled": [], "bare": True},
    "unitree_h1/configs": {
        "standin": "h1_mj_description", "disabled": _H1_DISABLED,
        "robot": {"base_link": "universe"},
        "group": {"left_arm": {"joints": ["left_shoulder_pitch", "left_shoulder_roll", "left_shoulder_yaw", "left_elbow"],
                               "named_states": {"home": [0.0] * 4}},
                  "right_arm": {"joints": ["right_shoulder_pitch", "right_shoulder_roll", "right_shoulder_yaw", "right_elbow"],
                                "named_states": {"home": [0.0] * 4}}},
        "acceleration_limits": {f"{s}_{j}": 1.0 for s in ("left", "right")
                                for j in ("shoulder_pitch", "shoulder_roll", "shoulder_yaw", "elbow")},
    },
}

# "boxy" variants: same kinematics, groups and disabled pairs, non-capsule link hulls (shaped_hull)
CONFIGS["panda_boxy/configs"] = {**CONFIGS["panda/configs"], "standin": "panda_boxy_description"}
CONFIGS["unitree_h1_boxy/configs"] = {**CONFIGS["unitree_h1/configs"], "standin": "h1_boxy_description"}
