"""Robot-description readers: a URDF reader, the xacro subset RAMP's files use, an MJCF
subset reader, and the SRDF ``disable_collisions`` reader.

Setup-time host code: it turns a description into the ``Model`` /
``GeometryModel`` containers of ``ramp_b200.kinematics`` following Pinocchio's URDF/MJCF
conventions as restated in SURVEY App. C.1, so that the reference's in-tree robots
(reference robots/rrr/rrr.urdf.xacro, robots/planar_rrr/robot.urdf.xacro, robots/rrr_mj/rrr.xml,
robots/acrobot/acrobot.xml, robots/floating.urdf, robots/planar.urdf,
tests/double_pendulum.urdf) load without any third-party wheel.  It replaces
``pinocchio.buildModelsFromUrdf`` / ``buildModelsFromMJCF`` / ``xacrodoc`` at
reference src/ramp/pinocchio_utils.py:209-286.  Collision meshes (URDF ``<mesh>``, MJCF
``type="mesh"``) become the convex hull of the file's vertices (ramp_b200.meshes; STL / OBJ,
``package://`` resolved against the description's directory as the reference's
``package_dirs`` does); a collision element that cannot be turned into a shape raises
``RobotDescriptionError`` -- none is ever skipped, a robot with a missing body would answer
"no collision" where the reference answers "collision".
"""

from __future__ import annotations

import math
import re
import xml.etree.ElementTree as ET
from pathlib import Path

import numpy as np

from ramp_b200 import meshes, shapes
from ramp_b200.exceptions import RobotDescriptionError
from ramp_b200.kinematics import (
    BODY,
    FIXED_JOINT,
    JOINT,
    Frame,
    GeometryModel,
    GeometryObject,
    Model,
)
from ramp_b200.se3 import SE3, Quaternion, rpy_to_matrix

XACRO_NS = "{http://www.ros.org/wiki/xacro}"


# --------------------------------------------------------------------------- xacro
def _xacro_eval(text: str, args: dict, props: dict) -> str:
    if text is None:
        return text

    def arg_sub(m):
        name = m.group(1)
        if name not in args:
            raise KeyError(f"xacro arg '{name}' is not defined")
        return str(args[name])

    text = re.sub(r"\$\(arg\s+([^)]+?)\s*\)", arg_sub, text)

    def expr_sub(m):
        env = {"pi": math.pi, "sin": math.sin, "cos": math.cos, "sqrt": math.sqrt,
               "radians": math.radians, "True": True, "False": False, **props}
        val = eval(m.group(1), {"__builtins__": {}}, env)  # noqa: S307 - trusted local files
        return str(val)

    return re.sub(r"\$\{([^}]*)\}", expr_sub, text)


def _truthy(s: str) -> bool:
    s = s.strip().lower()
    if s in ("true", "1", "1.0"):
        return True
    if s in ("false", "0", "0.0", ""):
        return False
    raise ValueError(f"xacro condition '{s}' is not a boolean")


def _coerce(value: str):
    try:
        return int(value)
    except ValueError:
        pass
    try:
        return float(value)
    except ValueError:
        return value


def render_xacro(source: "str | Path", mappings: dict | None = None) -> ET.Element:
    """Expand xacro:arg / xacro:property / $(arg) / ${expr} / xacro:if / xacro:unless and
    parameter-less xacro:macro definitions+calls.  ``source`` is a Path (file) or an XML
    string.  Returns the expanded root element."""
    root = ET.parse(source).getroot() if isinstance(source, Path) else ET.fromstring(source)
    args = dict(mappings or {})
    props: dict = {}
    macros: dict[str, ET.Element] = {}

    def expand(parent_out: ET.Element, elem: ET.Element):
        tag = elem.tag
        if tag.startswith(XACRO_NS):
            name = tag[len(XACRO_NS):]
            if name == "arg":
                args.setdefault(elem.get("name"), _xacro_eval(elem.get("default", ""), args, props))
            elif name == "property":
                props[elem.get("name")] = _coerce(_xacro_eval(elem.get("value", ""), args, props))
            elif name in ("if", "unless"):
                cond = _truthy(_xacro_eval(elem.get("value"), args, props))
                if cond == (name == "if"):
                    for child in elem:
                        expand(parent_out, child)
            elif name == "macro":
                macros[elem.get("name")] = elem
            elif name in macros:
                for child in macros[name]:
                    expand(parent_out, child)
            else:
                raise ValueError(f"unsupported xacro tag '{name}'")
            return
        out = ET.SubElement(parent_out, tag)
        for k, v in elem.attrib.items():
            out.set(k, _xacro_eval(v, args, props))
        if elem.text and elem.text.strip():
            out.text = _xacro_eval(elem.text, args, props)
        for child in elem:
            expand(out, child)

    holder = ET.Element("holder")
    expand(holder, root)
    return holder[0]


# ---------------------------------------------------------------------------- URDF
def _floats(s: str | None, n: int, default=0.0) -> np.ndarray:
    if s is None:
        return np.full(n, default, dtype=np.float64)
    v = np.array([float(x) for x in s.split()], dtype=np.float64)
    assert len(v) == n, f"expected {n} numbers, got '{s}'"
    return v


def _origin(elem: ET.Element | None) -> SE3:
    if elem is None:
        return SE3.Identity()
    xyz = _floats(elem.get("xyz"), 3)
    rpy = _floats(elem.get("rpy"), 3)
    return SE3(rpy_to_matrix(*rpy), xyz)


def _aligned_suffix(axis: np.ndarray) -> str | None:
    for k, s in enumerate("XYZ"):
        e = np.zeros(3)
        e[k] = 1.0
        if np.array_equal(axis, e):
            return s
    return None


def _urdf_geometry(geom: ET.Element, link_name: str, base_dir, package_dirs):
    if geom is None or len(geom) == 0:
        raise RobotDescriptionError(f"link '{link_name}': <collision> without a <geometry> shape")
    child = next(iter(geom))
    if child.tag == "box":
        return shapes.Box(_floats(child.get("size"), 3))
    if child.tag == "cylinder":
        return shapes.Cylinder(float(child.get("radius")), float(child.get("length")))
    if child.tag == "sphere":
        return shapes.Sphere(float(child.get("radius")))
    if child.tag == "capsule":
        return shapes.Capsule(float(child.get("radius")), float(child.get("length")))
    if child.tag == "mesh":
        filename = child.get("filename")
        if not filename:
            raise RobotDescriptionError(f"link '{link_name}': <mesh> without a filename")
        path = meshes.resolve_mesh_path(filename, base_dir, package_dirs)
        return meshes.load_convex(path, _floats(child.get("scale"), 3, 1.0))
    raise RobotDescriptionError(f"link '{link_name}': unknown collision geometry <{child.tag}>")


def build_models_from_urdf(root: ET.Element, base_dir=None, package_dirs=()) -> tuple[Model, GeometryModel, list[dict]]:
    """URDF element tree -> (Model, collision GeometryModel, mimic records).  ``base_dir`` resolves
    relative mesh paths, ``package_dirs`` the ``package://`` ones (the reference passes the
    description's own directory, src/ramp/pinocchio_utils.py:240-243).

    Pinocchio conventions restated: depth-first from the root link with
    children ordered by joint name (urdfdom keeps joints in a name-sorted map); no root joint
    is added; ``floating``/``planar`` joint TYPES become FreeFlyer/Planar joints; an axis that
    is exactly +x/+y/+z gives the aligned joint model, anything else the ``...Unaligned`` one;
    ``continuous`` gives the unbounded (cos, sin) model; fixed joints add a FIXED_JOINT frame
    and fold the child link into the parent joint; mimic tags are ignored by the kinematic
    model (RAMP enforces them itself).  Geometry objects are named ``<link>_<k>``.
    """
    assert root.tag == "robot", f"not a URDF: <{root.tag}>"
    model = Model(root.get("name", "robot"))
    links = {l.get("name"): l for l in root.findall("link")}
    joints = sorted(root.findall("joint"), key=lambda j: j.get("name"))
    children: dict[str, list[ET.Element]] = {}
    child_links = set()
    for j in joints:
        children.setdefault(j.find("parent").get("link"), []).append(j)
        child_links.add(j.find("child").get("link"))
    roots = [n for n in links if n not in child_links]
    assert len(roots) == 1, f"URDF must have exactly one root link, found {roots}"
    geom_model = GeometryModel()
    mimics: list[dict] = []

    def add_link_geoms(link_name: str, frame_id: int):
        frame = model.frames[frame_id]
        for k, col in enumerate(links[link_name].findall("collision")):
            geometry = _urdf_geometry(col.find("geometry"), link_name, base_dir, package_dirs)
            geom_model.addGeometryObject(
                GeometryObject(f"{link_name}_{k}", frame.parentJoint, frame_id,
                               frame.placement * _origin(col.find("origin")), geometry))

    def visit(link_name: str, frame_id: int):
        add_link_geoms(link_name, frame_id)
        parent_frame = model.frames[frame_id]
        for j in children.get(link_name, []):
            jname, jtype = j.get("name"), j.get("type")
            placement = parent_frame.placement * _origin(j.find("origin"))
            axis_el = j.find("axis")
            axis = _floats(axis_el.get("xyz") if axis_el is not None else None, 3)
            if axis_el is None:
                axis = np.array([1.0, 0.0, 0.0])
            limit = j.find("limit")
            child = j.find("child").get("link")
            if jtype == "fixed":
                fid = model.addFrame(Frame(jname, parent_frame.parentJoint, frame_id, placement, FIXED_JOINT))
                bid = model.addFrame(Frame(child, parent_frame.parentJoint, fid, placement, BODY))
                visit(child, bid)
                continue
            lower = upper = None
            suffix = _aligned_suffix(axis)
            if jtype in ("revolute", "prismatic"):
                base = "R" if jtype == "revolute" else "P"
                short = f"JointModel{base}{suffix}" if suffix else (
                    "JointModelRevoluteUnaligned" if jtype == "revolute" else "JointModelPrismaticUnaligned")
                if limit is not None:
                    lower, upper = float(limit.get("lower", 0.0)), float(limit.get("upper", 0.0))
            elif jtype == "continuous":
                short = f"JointModelRUB{suffix}" if suffix else "JointModelRevoluteUnboundedUnaligned"
            elif jtype == "floating":
                short = "JointModelFreeFlyer"
            elif jtype == "planar":
                short = "JointModelPlanar"
            else:
                raise ValueError(f"Unknown joint type: '{jtype}' for joint '{jname}'")
            jid = model.addJoint(parent_frame.parentJoint, short, placement, jname,
                                 axis=axis / np.linalg.norm(axis), lower=lower, upper=upper)
            fid = model.addFrame(Frame(jname, jid, frame_id, SE3.Identity(), JOINT))
            bid = model.addFrame(Frame(child, jid, fid, SE3.Identity(), BODY))
            mimic = j.find("mimic")
            if mimic is not None:
                mimics.append({"joint": jname, "mimicked": mimic.get("joint"),
                               "multiplier": float(mimic.get("multiplier", 1.0)),
                               "offset": float(mimic.get("offset", 0.0))})
            visit(child, bid)

    root_frame = model.addFrame(Frame(roots[0], 0, 0, SE3.Identity(), BODY))
    visit(roots[0], root_frame)
    return model, geom_model, mimics


# ---------------------------------------------------------------------------- MJCF
def _mj_quat(body: ET.Element) -> np.ndarray:
    if body.get("quat") is not None:
        w, x, y, z = _floats(body.get("quat"), 4)
        return Quaternion(w, x, y, z).normalize().matrix()
    if body.get("euler") is not None:
        e = _floats(body.get("euler"), 3)
        return rpy_to_matrix(e[0], e[1], e[2])
    return np.eye(3)


def build_models_from_mjcf(root: ET.Element, base_dir=None) -> tuple[Model, GeometryModel, list[dict]]:
    """MJCF subset (bodies, hinge/slide/free joints, sphere/box/cylinder/capsule/ellipsoid/mesh
    geoms, joint default classes for ``range``, ``<asset><mesh>`` with ``<compiler meshdir>``).
    Conventions restated from Pinocchio's MJCF parser:
    geom ``size`` holds half extents (doubled here), a body's frame sits at ``-joint.pos``
    in its joint frame, geoms with contype = conaffinity = 0 are visual only, geometry
    objects are named after the geom (``<body>Geom_<k>`` when unnamed)."""
    assert root.tag == "mujoco", f"not an MJCF: <{root.tag}>"
    model = Model(root.get("model", "mujoco"))
    geom_model = GeometryModel()
    compiler = root.find("compiler")
    meshdir = compiler.get("meshdir", "") if compiler is not None else ""
    # default classes: attributes of <joint> / <geom> / <mesh> per class, children inherit from
    # the enclosing <default>; the class-less top-level <default> is MuJoCo's "main"
    defaults: dict[str, dict[str, dict[str, str]]] = {"main": {}}

    def read_defaults(d: ET.Element, inherited: dict):
        mine = {k: dict(v) for k, v in inherited.items()}
        for tag in ("joint", "geom", "mesh"):
            el = d.find(tag)
            if el is not None:
                mine.setdefault(tag, {}).update(el.attrib)
        defaults[d.get("class") or "main"] = mine
        for child in d.findall("default"):
            read_defaults(child, mine)

    for d in root.findall("default"):
        read_defaults(d, {})

    def attrs(el: ET.Element, childclass: str | None) -> dict[str, str]:
        """Effective attributes of a joint / geom / mesh element: class defaults, then its own."""
        cls = el.get("class") or childclass or "main"
        if cls not in defaults:
            raise RobotDescriptionError(f"<{el.tag}> refers to default class '{cls}', which is not defined")
        return {**defaults[cls].get(el.tag, {}), **el.attrib}

    mesh_assets: dict[str, tuple[str, np.ndarray]] = {}
    for asset in root.findall("asset"):
        for me in asset.findall("mesh"):
            ma = attrs(me, None)
            file = ma.get("file")
            if file is None:
                continue  # inline vertex data: rejected where a collision geom asks for it
            mesh_assets[ma.get("name") or Path(file).stem] = (file, _floats(ma.get("scale"), 3, 1.0))

    def visit(body: ET.Element, parent_joint: int, parent_frame: int, parent_placement: SE3, childclass: str | None = None):
        name = body.get("name")
        childclass = body.get("childclass") or childclass
        pose = parent_placement * SE3(_mj_quat(body), _floats(body.get("pos"), 3))
        joint_id, frame_placement = parent_joint, pose
        jel = body.find("joint")
        if body.find("freejoint") is not None:
            fj = body.find("freejoint")
            joint_id = model.addJoint(parent_joint, "JointModelFreeFlyer", pose, fj.get("name", name))
            jf = model.addFrame(Frame(fj.get("name", name), joint_id, parent_frame, SE3.Identity(), JOINT))
            parent_frame, frame_placement = jf, SE3.Identity()
        elif jel is not None:
            ja = attrs(jel, childclass)
            jtype = ja.get("type", "hinge")
            jpos = _floats(ja.get("pos"), 3)
            axis = _floats(ja.get("axis"), 3, 0.0) if ja.get("axis") else np.array([0.0, 0.0, 1.0])
            suffix = _aligned_suffix(axis)
            if jtype == "hinge":
                short = f"JointModelR{suffix}" if suffix else "JointModelRevoluteUnaligned"
            elif jtype == "slide":
                short = f"JointModelP{suffix}" if suffix else "JointModelPrismaticUnaligned"
            elif jtype == "free":
                short = "JointModelFreeFlyer"
            else:
                raise ValueError(f"Unknown joint type: '{jtype}' for joint '{jel.get('name')}'")
            rng = tuple(_floats(ja.get("range"), 2)) if ja.get("range") else None
            joint_id = model.addJoint(parent_joint, short, pose * SE3(np.eye(3), jpos), jel.get("name"),
                                      axis=axis / np.linalg.norm(axis),
                                      lower=None if rng is None else rng[0],
                                      upper=None if rng is None else rng[1])
            jf = model.addFrame(Frame(jel.get("name"), joint_id, parent_frame, SE3.Identity(), JOINT))
            parent_frame, frame_placement = jf, SE3(np.eye(3), -jpos)
        bid = model.addFrame(Frame(name, joint_id, parent_frame, frame_placement, BODY))
        for k, g_el in enumerate(body.findall("geom")):
            g = attrs(g_el, childclass)
            if g.get("contype") == "0" and g.get("conaffinity") == "0":
                continue
            gtype = g.get("type", "sphere")
            size = [float(x) for x in g.get("size", "0").split()]
            fromto_pose = None
            if g.get("fromto") and gtype in ("capsule", "cylinder"):
                # the segment between two points: local z along it, half length from its extent
                ft = _floats(g.get("fromto"), 6)
                a, b = ft[:3], ft[3:]
                ez = b - a
                half = 0.5 * float(np.linalg.norm(ez))
                ez = ez / (2.0 * half) if half > 0 else np.array([0.0, 0.0, 1.0])
                ex = np.cross(ez, [1.0, 0.0, 0.0] if abs(ez[0]) < 0.9 else [0.0, 1.0, 0.0])
                ex /= np.linalg.norm(ex)
                fromto_pose = SE3(np.column_stack([ex, np.cross(ez, ex), ez]), 0.5 * (a + b))
                size = [size[0], half]
            if gtype == "sphere":
                geometry = shapes.Sphere(size[0])
            elif gtype == "box":
                geometry = shapes.Box(2.0 * np.asarray(size[:3]))
            elif gtype == "cylinder":
                geometry = shapes.Cylinder(size[0], 2.0 * size[1])
            elif gtype == "capsule":
                geometry = shapes.Capsule(size[0], 2.0 * size[1])
            elif gtype == "ellipsoid":
                geometry = shapes.Ellipsoid(np.asarray(size[:3]))
            elif gtype == "mesh":
                mname = g.get("mesh")
                if mname not in mesh_assets:
                    raise RobotDescriptionError(f"body '{name}': geom refers to mesh '{mname}', which <asset> does not define by file")
                file, scale = mesh_assets[mname]
                mpath = Path(file)
                if not mpath.is_absolute():
                    mpath = Path(base_dir or ".") / meshdir / mpath
                geometry = meshes.load_convex(mpath, scale)
            else:
                # plane / hfield / sdf: no finite convex stand-in exists; dropping the body would
                # turn "collision" answers into "free"
                raise RobotDescriptionError(f"body '{name}': collision geom of type '{gtype}' is not supported")
            gpl = frame_placement * (fromto_pose if fromto_pose is not None else SE3(_mj_quat(g), _floats(g.get("pos"), 3)))
            geom_model.addGeometryObject(
                GeometryObject(g.get("name") or f"{name}Geom_{k}", joint_id, bid, gpl, geometry))
        for child in body.findall("body"):
            visit(child, joint_id, bid, frame_placement, childclass)

    world = root.find("worldbody")
    for body in world.findall("body"):
        visit(body, 0, 0, SE3.Identity())
    return model, geom_model, []


# ---------------------------------------------------------------------------- SRDF
def read_disabled_collisions(root: ET.Element) -> list[tuple[str, str]]:
    return [(e.get("link1"), e.get("link2")) for e in root.iter("disable_collisions")]


def load_description(path: Path, mappings: dict | None = None):
    """Dispatch on the suffix like reference pinocchio_utils.load_models (:247-273)."""
    path = Path(path)
    if path.suffix in (".xacro", ".urdf"):
        base = path.parent.resolve()
        return build_models_from_urdf(render_xacro(path, mappings), base, [base])
    if path.suffix == ".xml":
        return build_models_from_mjcf(ET.parse(path).getroot(), path.parent.resolve())
    msg = f"Unknown robot description file type: {path}"
    raise ValueError(msg)
