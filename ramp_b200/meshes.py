"""Mesh files -> convex collision geometry (setup time, host only).

The reference gets file meshes through Pinocchio's URDF / MJCF loaders with ``package_dirs``
(reference src/ramp/pinocchio_utils.py:238-244, :264-270) and through ``hppfcl.MeshLoader``
(reference src/ramp/robot_model.py:429-435); coal then treats a mesh as a triangle soup
(``BVHModel``).  This build's narrow phase works on convex cores (GJK), so a mesh becomes the
CONVEX HULL of its vertices -- the stated deviation of SURVEY section 7 ("mesh semantics"), applied
to the oracle and to the GPU alike because both receive the same vertex array.

Formats: binary STL, ASCII STL, Wavefront OBJ (vertex positions only; faces are irrelevant to a
hull).  Anything else raises ``RobotDescriptionError`` -- a collision mesh is never skipped.

A hull with more vertices than ``max_vertices`` is reduced to a subset picked by farthest-point
sampling and then scaled about the hull's centroid by the smallest factor that makes the reduced
hull contain every original vertex again: the result is a conservative (never smaller)
approximation with at most ``max_vertices`` vertices.
"""

from __future__ import annotations

import struct
from pathlib import Path

import numpy as np

from ramp_b200.exceptions import RobotDescriptionError
from ramp_b200.shapes import Convex

MAX_HULL_VERTICES = 128


def _read_stl(data: bytes, path) -> np.ndarray:
    # binary: 80-byte header, uint32 triangle count, 50 bytes per triangle.  The size test comes
    # first: plenty of binary files start with the word "solid".
    if len(data) >= 84:
        (ntri,) = struct.unpack_from("<I", data, 80)
        if 84 + 50 * ntri == len(data):
            if ntri == 0:
                raise RobotDescriptionError(f"{path}: STL without triangles")
            rec = np.frombuffer(data, dtype=np.dtype([("n", "<f4", 3), ("v", "<f4", (3, 3)), ("a", "<u2")]), count=ntri, offset=84)
            return rec["v"].reshape(-1, 3).astype(np.float64)
    text = data.decode("ascii", errors="replace")
    if not text.lstrip().lower().startswith("solid"):
        raise RobotDescriptionError(f"{path}: neither a binary nor an ASCII STL")
    pts = [ln.split()[1:4] for ln in text.splitlines() if ln.strip().lower().startswith("vertex")]
    if not pts:
        raise RobotDescriptionError(f"{path}: ASCII STL without vertices")
    return np.asarray(pts, dtype=np.float64)


def _read_obj(data: bytes, path) -> np.ndarray:
    pts = []
    for ln in data.decode("utf-8", errors="replace").splitlines():
        f = ln.split()
        if len(f) >= 4 and f[0] == "v":
            pts.append(f[1:4])
    if not pts:
        raise RobotDescriptionError(f"{path}: OBJ without vertices")
    return np.asarray(pts, dtype=np.float64)


def read_mesh_vertices(path) -> np.ndarray:
    """float64[V, 3] vertex positions of a mesh file (duplicates included)."""
    path = Path(path)
    if not path.exists():
        raise RobotDescriptionError(f"mesh file not found: {path}")
    suffix = path.suffix.lower()
    data = path.read_bytes()
    if suffix == ".stl":
        pts = _read_stl(data, path)
    elif suffix == ".obj":
        pts = _read_obj(data, path)
    else:
        raise RobotDescriptionError(
            f"{path}: unsupported collision mesh format '{suffix}' (STL and OBJ are read); "
            "convert the mesh or replace it with a primitive -- it is not skipped")
    if not np.isfinite(pts).all():
        raise RobotDescriptionError(f"{path}: non-finite vertex coordinates")
    return pts


def _hull_indices(points: np.ndarray) -> np.ndarray:
    """Indices of the points on the convex hull; degenerate (flat / collinear) sets keep all
    their distinct points, which is a valid vertex list for GJK."""
    if len(points) < 5:
        return np.arange(len(points))
    try:
        from scipy.spatial import ConvexHull, QhullError
    except ImportError as e:  # pragma: no cover - scipy ships with the image
        raise RobotDescriptionError("scipy is needed to build convex hulls of mesh files") from e
    try:
        return np.sort(ConvexHull(points).vertices)
    except QhullError:
        return np.arange(len(points))


def _facets(points: np.ndarray):
    from scipy.spatial import ConvexHull

    eq = ConvexHull(points).equations  # n . x + d <= 0 inside
    return eq[:, :3], -eq[:, 3]


def convex_hull_vertices(points, max_vertices: int = MAX_HULL_VERTICES) -> np.ndarray:
    """Vertices of the convex hull of ``points`` (float64[V, 3], V <= max_vertices)."""
    pts = np.unique(np.round(np.asarray(points, dtype=np.float64).reshape(-1, 3), 12), axis=0)
    hull = pts[_hull_indices(pts)]
    if len(hull) <= max_vertices:
        return hull
    # farthest-point sampling on the hull vertices, seeded with the extreme points of the axes
    c = hull.mean(axis=0)
    chosen = list(dict.fromkeys([int(np.argmin(hull[:, k])) for k in range(3)] + [int(np.argmax(hull[:, k])) for k in range(3)]))
    d = np.min(np.linalg.norm(hull[:, None, :] - hull[chosen][None, :, :], axis=2), axis=1)
    while len(chosen) < max_vertices:
        i = int(np.argmax(d))
        chosen.append(i)
        d = np.minimum(d, np.linalg.norm(hull - hull[i], axis=1))
    sub = hull[chosen]
    # scale about the centroid until every dropped vertex is inside again
    try:
        n, off = _facets(sub)
    except Exception:  # flat subset: keep it as it is
        return sub
    num = (hull - c) @ n.T              # [V, F] signed offsets of the vertices from c along the facet normals
    den = off - n @ c                   # [F] facet offsets from c (> 0: c is interior)
    ok = den > 1e-12
    scale = max(1.0, float((num[:, ok] / den[ok]).max()))
    return c + (sub - c) * (scale * (1.0 + 1e-9))


def load_convex(path, scale=(1.0, 1.0, 1.0), max_vertices: int = MAX_HULL_VERTICES) -> Convex:
    """The convex hull of a mesh file as a ``Convex`` shape; ``scale`` is the per-axis factor of a
    URDF ``<mesh scale=...>`` / MJCF ``<mesh scale=...>`` element."""
    pts = read_mesh_vertices(path) * np.asarray(scale, dtype=np.float64).reshape(3)
    return Convex(convex_hull_vertices(pts, max_vertices))


def resolve_mesh_path(filename: str, base_dir: Path | None, package_dirs=()) -> Path:
    """``package://<pkg>/<rel>`` is looked up as ``<dir>/<pkg>/<rel>`` and ``<dir>/<rel>`` for
    every directory of ``package_dirs`` (Pinocchio's rule for ``package_dirs``, which the
    reference sets to the description's own directory, src/ramp/pinocchio_utils.py:240-243);
    ``file://`` and plain paths are taken as they are, relative ones against ``base_dir``."""
    if filename.startswith("package://"):
        rest = filename[len("package://"):]
        pkg, _, rel = rest.partition("/")
        tried = []
        for d in package_dirs:
            for cand in (Path(d) / pkg / rel, Path(d) / rel, Path(d).parent / pkg / rel):
                tried.append(str(cand))
                if cand.exists():
                    return cand
        raise RobotDescriptionError(f"cannot resolve '{filename}'; tried {tried}")
    if filename.startswith("file://"):
        filename = filename[len("file://"):]
    p = Path(filename)
    if not p.is_absolute() and base_dir is not None:
        p = Path(base_dir) / p
    return p
