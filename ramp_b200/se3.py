"""Minimal SE(3) / quaternion value types with the Pinocchio surface RAMP uses.

RAMP's call sites build poses as ``pinocchio.SE3(pinocchio.Quaternion(w, x, y, z), t)``
(reference tests/motion_planner_test.py:36-40, examples/motion_planning.py:56-60) and read
``.np``, ``.translation`` and ``.rotation`` off the result of ``get_frame_pose``
(reference src/ramp/motion_planner.py:497-499).  These classes keep that surface so the
same user code runs against the GPU path.
"""

from __future__ import annotations

import numpy as np


class Quaternion:
    """Eigen-style quaternion: constructor order is (w, x, y, z)."""

    __slots__ = ("w", "x", "y", "z")

    def __init__(self, w=1.0, x=0.0, y=0.0, z=0.0):
        if isinstance(w, np.ndarray) and w.shape == (3, 3):
            q = Quaternion.from_matrix(w)
            w, x, y, z = q.w, q.x, q.y, q.z
        self.w, self.x, self.y, self.z = float(w), float(x), float(y), float(z)

    @staticmethod
    def Identity():  # noqa: N802  (Pinocchio spelling)
        return Quaternion(1.0, 0.0, 0.0, 0.0)

    @staticmethod
    def from_matrix(r: np.ndarray) -> "Quaternion":
        r = np.asarray(r, dtype=np.float64)
        t = np.trace(r)
        if t > 0:
            s = np.sqrt(t + 1.0) * 2
            return Quaternion(0.25 * s, (r[2, 1] - r[1, 2]) / s, (r[0, 2] - r[2, 0]) / s, (r[1, 0] - r[0, 1]) / s)
        i = int(np.argmax(np.diag(r)))
        j, k = (i + 1) % 3, (i + 2) % 3
        s = np.sqrt(r[i, i] - r[j, j] - r[k, k] + 1.0) * 2
        v = [0.0, 0.0, 0.0]
        v[i] = 0.25 * s
        v[j] = (r[j, i] + r[i, j]) / s
        v[k] = (r[k, i] + r[i, k]) / s
        return Quaternion((r[k, j] - r[j, k]) / s, *v)

    def norm(self) -> float:
        return float(np.sqrt(self.w**2 + self.x**2 + self.y**2 + self.z**2))

    def normalize(self) -> "Quaternion":
        n = self.norm()
        self.w, self.x, self.y, self.z = self.w / n, self.x / n, self.y / n, self.z / n
        return self

    def coeffs(self) -> np.ndarray:
        """Eigen coefficient order (x, y, z, w)."""
        return np.array([self.x, self.y, self.z, self.w])

    def matrix(self) -> np.ndarray:
        """Eigen ``toRotationMatrix``: assumes unit norm and does NOT normalise.

        The reference passes ``Quaternion(0.707, 0.707, 0, 0)`` un-normalised; the resulting
        (slightly non-orthonormal) matrix is what Pinocchio and coal then use, so it is
        what the oracle and the GPU use too.
        """
        x, y, z, w = self.x, self.y, self.z, self.w
        tx, ty, tz = 2 * x, 2 * y, 2 * z
        twx, twy, twz = tx * w, ty * w, tz * w
        txx, txy, txz = tx * x, ty * x, tz * x
        tyy, tyz, tzz = ty * y, tz * y, tz * z
        return np.array(
            [
                [1 - (tyy + tzz), txy - twz, txz + twy],
                [txy + twz, 1 - (txx + tzz), tyz - twx],
                [txz - twy, tyz + twx, 1 - (txx + tyy)],
            ]
        )

    toRotationMatrix = matrix  # noqa: N815

    def __repr__(self):
        return f"Quaternion(w={self.w}, x={self.x}, y={self.y}, z={self.z})"


class SE3:
    """Rigid transform (R, t); ``a * b`` composes as (Ra Rb, ta + Ra tb)."""

    __slots__ = ("rotation", "translation")

    def __init__(self, rotation=None, translation=None):
        if rotation is None:
            rotation = np.eye(3)
        if isinstance(rotation, Quaternion):
            rotation = rotation.matrix()
        rotation = np.asarray(rotation, dtype=np.float64)
        if rotation.shape == (4, 4) and translation is None:
            translation = rotation[:3, 3]
            rotation = rotation[:3, :3]
        if translation is None:
            translation = np.zeros(3)
        self.rotation = np.array(rotation, dtype=np.float64).reshape(3, 3)
        self.translation = np.array(translation, dtype=np.float64).reshape(3)

    @staticmethod
    def Identity():  # noqa: N802
        return SE3()

    @staticmethod
    def from_rt34(m) -> "SE3":
        m = np.asarray(m, dtype=np.float64).reshape(3, 4)
        return SE3(m[:, :3], m[:, 3])

    def rt34(self) -> np.ndarray:
        """3x4 row-major [R|t] (the layout of every table and kernel in this repo)."""
        return np.concatenate([self.rotation, self.translation[:, None]], axis=1)

    @property
    def np(self) -> "np.ndarray":
        h = np.eye(4)
        h[:3, :3] = self.rotation
        h[:3, 3] = self.translation
        return h

    homogeneous = np

    def __mul__(self, other: "SE3") -> "SE3":
        return SE3(self.rotation @ other.rotation, self.translation + self.rotation @ other.translation)

    def act(self, p):
        return self.rotation @ np.asarray(p, dtype=np.float64) + self.translation

    def inverse(self) -> "SE3":
        return SE3(self.rotation.T, -self.rotation.T @ self.translation)

    def copy(self) -> "SE3":
        return SE3(self.rotation.copy(), self.translation.copy())

    def __repr__(self):
        return f"SE3(R=\n{self.rotation},\n t={self.translation})"


def rpy_to_matrix(roll: float, pitch: float, yaw: float) -> np.ndarray:
    """URDF rpy convention: R = Rz(yaw) Ry(pitch) Rx(roll)."""
    cr, sr = np.cos(roll), np.sin(roll)
    cp, sp = np.cos(pitch), np.sin(pitch)
    cy, sy = np.cos(yaw), np.sin(yaw)
    return np.array(
        [
            [cy * cp, cy * sp * sr - sy * cr, cy * sp * cr + sy * sr],
            [sy * cp, sy * sp * sr + cy * cr, sy * sp * cr - cy * sr],
            [-sp, cp * sr, cp * cr],
        ]
    )


def xyzquat_to_se3(pose) -> SE3:
    """pinocchio.XYZQUATToSE3: [x, y, z, qx, qy, qz, qw]."""
    p = np.asarray(pose, dtype=np.float64)
    return SE3(Quaternion(p[6], p[3], p[4], p[5]), p[:3])
