"""Collision meshes: STL / OBJ readers, convex hulls, URDF <mesh> and MJCF type="mesh" elements.

The reference loads file meshes through Pinocchio's loaders (reference
src/ramp/pinocchio_utils.py:238-244, :264-270); here a mesh becomes the convex hull of its
vertices and an element that cannot be read raises instead of being skipped.  CPU tests: readers,
hull builder, description wiring, and the oracle's verdicts on a mesh cube against the same cube
as a ``Box``.  The GPU half (same verdicts on the device) is in test_gpu_parity.py.
"""

import struct
import xml.etree.ElementTree as ET
from pathlib import Path

import numpy as np
import pytest

GOLDEN = Path(__file__).resolve().parent / "golden"

CUBE_TRIS = [(0, 2, 1), (0, 3, 2), (4, 5, 6), (4, 6, 7), (0, 1, 5), (0, 5, 4), (2, 3, 7), (2, 7, 6), (1, 2, 6), (1, 6, 5),
             (0, 4, 7), (0, 7, 3)]


def cube_corners(sx, sy, sz):
    return np.array([[-sx, -sy, -sz], [sx, -sy, -sz], [sx, sy, -sz], [-sx, sy, -sz],
                     [-sx, -sy, sz], [sx, -sy, sz], [sx, sy, sz], [-sx, sy, sz]], dtype=np.float64) * 0.5


def binary_stl(corners) -> bytes:
    out = bytearray(b"ramp-b200 test cube".ljust(80, b" "))
    out += struct.pack("<I", len(CUBE_TRIS))
    for tri in CUBE_TRIS:
        out += struct.pack("<3f", 0.0, 0.0, 0.0)
        for i in tri:
            out += struct.pack("<3f", *corners[i])
        out += struct.pack("<H", 0)
    return bytes(out)


def ascii_stl(corners) -> str:
    lines = ["solid cube"]
    for tri in CUBE_TRIS:
        lines += ["facet normal 0 0 0", " outer loop"] + [" vertex %.9g %.9g %.9g" % tuple(corners[i]) for i in tri] + [" endloop", "endfacet"]
    return "\n".join(lines + ["endsolid cube"])


def obj_text(corners) -> str:
    return "\n".join(["# cube"] + ["v %.9g %.9g %.9g" % tuple(c) for c in corners] + ["f %d %d %d" % tuple(i + 1 for i in t) for t in CUBE_TRIS])


def test_committed_fixture_is_the_12_triangle_cube():
    data = (GOLDEN / "unit_cube.stl").read_bytes()
    assert data == binary_stl(cube_corners(1, 1, 1)) and len(data) == 84 + 50 * 12


@pytest.mark.parametrize("kind", ["binary_stl", "ascii_stl", "obj"])
def test_readers_return_the_file_vertices(tmp_path, kind):
    from ramp_b200 import meshes

    c = cube_corners(0.2, 0.1, 0.3)
    if kind == "binary_stl":
        p = tmp_path / "cube.stl"
        p.write_bytes(binary_stl(c))
    elif kind == "ascii_stl":
        p = tmp_path / "cube.STL"
        p.write_text(ascii_stl(c))
    else:
        p = tmp_path / "cube.obj"
        p.write_text(obj_text(c))
    v = meshes.read_mesh_vertices(p)
    assert len(v) == (8 if kind == "obj" else 36)
    hull = meshes.convex_hull_vertices(v)
    assert len(hull) == 8
    assert np.allclose(np.sort(hull.view([("", hull.dtype)] * 3), axis=0).view(hull.dtype).reshape(-1, 3),
                       np.sort(c.view([("", c.dtype)] * 3), axis=0).view(c.dtype).reshape(-1, 3), atol=1e-6)


def test_unreadable_meshes_raise(tmp_path):
    from ramp_b200 import meshes
    from ramp_b200.exceptions import RobotDescriptionError

    with pytest.raises(RobotDescriptionError, match="not found"):
        meshes.read_mesh_vertices(tmp_path / "nope.stl")
    (tmp_path / "m.dae").write_text("<COLLADA/>")
    with pytest.raises(RobotDescriptionError, match="unsupported"):
        meshes.read_mesh_vertices(tmp_path / "m.dae")
    (tmp_path / "junk.stl").write_bytes(b"\x00" * 100)
    with pytest.raises(RobotDescriptionError):
        meshes.read_mesh_vertices(tmp_path / "junk.stl")
    (tmp_path / "empty.obj").write_text("# nothing\n")
    with pytest.raises(RobotDescriptionError):
        meshes.read_mesh_vertices(tmp_path / "empty.obj")


def test_large_hulls_are_reduced_conservatively():
    """More hull vertices than the cap: the reduced hull has at most 128 vertices and contains
    every vertex of the full one (never smaller than the mesh)."""
    from scipy.spatial import ConvexHull

    from ramp_b200 import meshes

    rng = np.random.default_rng(3)
    v = rng.normal(size=(3000, 3))
    v = v / np.linalg.norm(v, axis=1, keepdims=True) * np.array([0.3, 0.2, 0.1]) + np.array([0.05, -0.02, 0.4])
    full = meshes.convex_hull_vertices(v, max_vertices=10**6)
    assert len(full) > 1000
    red = meshes.convex_hull_vertices(v)
    assert len(red) <= meshes.MAX_HULL_VERTICES
    eq = ConvexHull(red).equations
    assert (full @ eq[:, :3].T + eq[:, 3]).max() <= 1e-9
    # and it is tight: inflated by a few per cent at most
    assert np.abs(red - red.mean(0)).max() <= 1.08 * np.abs(full - full.mean(0)).max()


URDF = """<robot name="meshy">
  <link name="base"><collision><origin xyz="0 0 0.1" rpy="0 0 0"/><geometry>{base}</geometry></collision></link>
  <link name="arm"><collision><origin xyz="0 0 0.25" rpy="0 0.1 0"/><geometry>{arm}</geometry></collision></link>
  <link name="tip"><collision><geometry><sphere radius="0.06"/></geometry></collision></link>
  <joint name="j1" type="revolute"><parent link="base"/><child link="arm"/><origin xyz="0 0 0.2" rpy="0 0 0"/>
    <axis xyz="0 1 0"/><limit lower="-3" upper="3" effort="1" velocity="1"/></joint>
  <joint name="j2" type="revolute"><parent link="arm"/><child link="tip"/><origin xyz="0 0 0.5" rpy="0 0 0"/>
    <axis xyz="1 0 0"/><limit lower="-3" upper="3" effort="1" velocity="1"/></joint>
</robot>"""


def _write_mesh_robot(tmp_path):
    """The same robot twice: boxes as URDF primitives, and as STL cubes scaled by <mesh scale>."""
    (tmp_path / "pkg" / "meshes").mkdir(parents=True)
    (tmp_path / "pkg" / "meshes" / "unit_cube.stl").write_bytes((GOLDEN / "unit_cube.stl").read_bytes())
    (tmp_path / "rel_cube.obj").write_text(obj_text(cube_corners(1, 1, 1)))
    prim = URDF.format(base='<box size="0.3 0.3 0.2"/>', arm='<box size="0.08 0.1 0.5"/>')
    mesh = URDF.format(base='<mesh filename="package://pkg/meshes/unit_cube.stl" scale="0.3 0.3 0.2"/>',
                       arm='<mesh filename="rel_cube.obj" scale="0.08 0.1 0.5"/>')
    (tmp_path / "prim.urdf").write_text(prim)
    (tmp_path / "mesh.urdf").write_text(mesh)
    return tmp_path / "prim.urdf", tmp_path / "mesh.urdf"


def test_urdf_mesh_elements_become_convex_hulls_with_the_box_verdicts(built, tmp_path):
    from oracle import oracle
    from ramp_b200 import workloads
    from ramp_b200.desc import ModelDesc
    from ramp_b200.robot_model import load_robot_model
    from ramp_b200.robot_state import RobotState
    from ramp_b200.shapes import Box, Convex

    prim_path, mesh_path = _write_mesh_robot(tmp_path)
    rm_p, rm_m = load_robot_model(prim_path), load_robot_model(mesh_path)
    gp, gm = rm_p.collision_model.geometryObjects, rm_m.collision_model.geometryObjects
    assert [g.name for g in gp] == [g.name for g in gm] and len(gm) == 3
    assert isinstance(gp[0].geometry, Box) and isinstance(gm[0].geometry, Convex) and isinstance(gm[1].geometry, Convex)
    assert len(gm[0].geometry.points) == 8 and np.allclose(np.abs(gm[0].geometry.points).max(0), [0.15, 0.15, 0.1])
    assert np.allclose(np.abs(gm[1].geometry.points).max(0), [0.04, 0.05, 0.25])
    for rm in (rm_p, rm_m):
        RobotState(rm).add_object(workloads.reference_capsule())
    q = np.random.default_rng(5).uniform(-3, 3, (4000, 2))
    a = oracle.check(ModelDesc(rm_p.model, rm_p.collision_model), q)
    b = oracle.check(ModelDesc(rm_m.model, rm_m.collision_model), q)
    outside = np.abs(a["pair_dist"]) > 1e-6
    assert ((a["pair_dist"] <= 0) == (b["pair_dist"] <= 0))[outside].all()
    assert np.abs(a["pair_dist"] - b["pair_dist"])[a["pair_dist"] > 0].max() < 1e-6
    ph = a["pair_dist"] <= 0
    assert (ph.any(axis=0) & ~ph.all(axis=0)).sum() >= 2  # several pairs go both ways over the sample


def test_urdf_collision_elements_are_never_dropped(tmp_path):
    from ramp_b200 import description
    from ramp_b200.exceptions import RobotDescriptionError

    for geom in ('<mesh filename="package://nowhere/x.stl"/>', '<mesh filename="visual.dae"/>', "<mesh/>", '<torus r="1"/>', ""):
        (tmp_path / "visual.dae").write_text("<COLLADA/>")
        root = ET.fromstring(f'<robot name="r"><link name="l"><collision><geometry>{geom}</geometry></collision></link></robot>')
        with pytest.raises(RobotDescriptionError):
            description.build_models_from_urdf(root, tmp_path, [tmp_path])


MJCF = """<mujoco model="m"><compiler meshdir="assets"/>
<default><default class="visual"><geom contype="0" conaffinity="0" group="2"/></default>
 <default class="collision"><geom group="3" type="mesh"/></default>
 <default class="arm"><joint range="-2 2" axis="0 1 0"/></default></default>
<asset><mesh name="cube" file="unit_cube.stl" scale="0.2 0.2 0.4"/><mesh file="other.obj"/></asset>
<worldbody><body name="b0" pos="0 0 0.3" childclass="arm"><joint name="j0"/>
 <geom class="visual" mesh="cube"/><geom class="collision" mesh="cube" pos="0 0 0.2"/>
 <geom type="capsule" size="0.03" fromto="0 0 0 0.2 0 0"/>
 <body name="b1" pos="0 0 0.5"><joint name="j1" axis="1 0 0"/><geom class="collision" mesh="other"/><geom type="ellipsoid" size="0.1 0.05 0.02"/></body>
</body></worldbody></mujoco>"""


def test_mjcf_mesh_geoms_default_classes_and_fromto(tmp_path):
    from ramp_b200 import description
    from ramp_b200.exceptions import RobotDescriptionError
    from ramp_b200.shapes import Capsule, Convex, Ellipsoid

    (tmp_path / "assets").mkdir()
    (tmp_path / "assets" / "unit_cube.stl").write_bytes((GOLDEN / "unit_cube.stl").read_bytes())
    (tmp_path / "assets" / "other.obj").write_text(obj_text(cube_corners(0.1, 0.1, 0.1)))
    model, gm, _ = description.build_models_from_mjcf(ET.fromstring(MJCF), tmp_path)
    kinds = [type(g.geometry) for g in gm.geometryObjects]
    assert kinds == [Convex, Capsule, Convex, Ellipsoid]  # the visual-class geom is not a collision body
    assert np.allclose(np.abs(gm.geometryObjects[0].geometry.points).max(0), [0.1, 0.1, 0.2])
    cap = gm.geometryObjects[1]
    assert np.isclose(cap.geometry.halfLength, 0.1) and np.allclose(cap.placement.rotation[:, 2], [1, 0, 0])
    assert [j.shortname() for j in model.joints[1:]] == ["JointModelRY", "JointModelRX"]
    assert np.allclose(model.lowerPositionLimit, [-2, -2]) and np.allclose(model.upperPositionLimit, [2, 2])
    bad = MJCF.replace('<geom type="ellipsoid" size="0.1 0.05 0.02"/>', '<geom type="plane" size="1 1 1"/>')
    with pytest.raises(RobotDescriptionError, match="plane"):
        description.build_models_from_mjcf(ET.fromstring(bad), tmp_path)
    with pytest.raises(RobotDescriptionError, match="mesh 'ghost'"):
        description.build_models_from_mjcf(ET.fromstring(MJCF.replace('mesh="other"', 'mesh="ghost"')), tmp_path)
