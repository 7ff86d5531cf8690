"""Collision shape value types with the coal/hpp-fcl constructor conventions.

RAMP builds scene objects from ``hppfcl.Capsule(radius, lz)``, ``hppfcl.Cylinder(radius, lz)``,
``hppfcl.Box(x, y, z)``, ``hppfcl.Sphere(radius)``, ``hppfcl.Ellipsoid(rx, ry, rz)`` (reference
src/ramp/robot_model.py:429-435, tests/motion_planner_test.py:36) and gets Box/Cylinder/Sphere
from URDF (reference robots/rrr/rrr.urdf.xacro:17-103).  ``lz`` and box sides are FULL lengths,
as in coal; the device tables store half lengths (include/ramp_b200.h RampShapeType).
"""

from __future__ import annotations

import numpy as np

SHAPE_SPHERE, SHAPE_CAPSULE, SHAPE_BOX, SHAPE_CYLINDER, SHAPE_CONVEX, SHAPE_ELLIPSOID = range(6)


class CollisionGeometry:
    shape_type: int = -1

    def params(self) -> tuple[float, float, float, float]:
        raise NotImplementedError

    def bounding_radius(self) -> float:
        """Radius of a sphere centred at the local origin that contains the shape."""
        raise NotImplementedError

    def aabb_half_extents(self) -> np.ndarray:
        """Half extents of the local-frame axis-aligned box that contains the shape."""
        raise NotImplementedError


class Sphere(CollisionGeometry):
    shape_type = SHAPE_SPHERE

    def __init__(self, radius: float):
        self.radius = float(radius)

    def params(self):
        return (self.radius, 0.0, 0.0, 0.0)

    def bounding_radius(self):
        return self.radius

    def aabb_half_extents(self):
        return np.full(3, self.radius)


class Capsule(CollisionGeometry):
    """Segment (0, 0, +-lz/2) swept by ``radius``."""

    shape_type = SHAPE_CAPSULE

    def __init__(self, radius: float, lz: float):
        self.radius, self.halfLength = float(radius), 0.5 * float(lz)  # noqa: N815

    def params(self):
        return (self.radius, self.halfLength, 0.0, 0.0)

    def bounding_radius(self):
        return self.radius + self.halfLength

    def aabb_half_extents(self):
        return np.array([self.radius, self.radius, self.radius + self.halfLength])


class Cylinder(CollisionGeometry):
    shape_type = SHAPE_CYLINDER

    def __init__(self, radius: float, lz: float):
        self.radius, self.halfLength = float(radius), 0.5 * float(lz)  # noqa: N815

    def params(self):
        return (self.radius, self.halfLength, 0.0, 0.0)

    def bounding_radius(self):
        return float(np.hypot(self.radius, self.halfLength))

    def aabb_half_extents(self):
        return np.array([self.radius, self.radius, self.halfLength])


class Box(CollisionGeometry):
    shape_type = SHAPE_BOX

    def __init__(self, x, y=None, z=None):
        sides = np.asarray(x if y is None else [x, y, z], dtype=np.float64).reshape(3)
        self.halfSide = 0.5 * sides  # noqa: N815

    def params(self):
        return (*self.halfSide.tolist(), 0.0)

    def bounding_radius(self):
        return float(np.linalg.norm(self.halfSide))

    def aabb_half_extents(self):
        return self.halfSide.copy()


class Ellipsoid(CollisionGeometry):
    shape_type = SHAPE_ELLIPSOID

    def __init__(self, x, y=None, z=None):
        self.radii = np.asarray(x if y is None else [x, y, z], dtype=np.float64).reshape(3)

    def params(self):
        return (*self.radii.tolist(), 0.0)

    def bounding_radius(self):
        return float(self.radii.max())

    def aabb_half_extents(self):
        return self.radii.copy()


class Convex(CollisionGeometry):
    """Convex hull given by its vertices (only the vertex set matters to GJK)."""

    shape_type = SHAPE_CONVEX

    def __init__(self, points):
        self.points = np.ascontiguousarray(points, dtype=np.float64).reshape(-1, 3)
        if len(self.points) == 0:
            raise ValueError("Convex needs at least one vertex")

    def params(self):
        return (0.0, 0.0, 0.0, 0.0)

    def bounding_radius(self):
        return float(np.linalg.norm(self.points, axis=1).max())

    def aabb_half_extents(self):
        return np.abs(self.points).max(axis=0)
