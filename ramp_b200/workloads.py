"""The BASELINE.json workloads as (robot model, scene, synthetic sampler) bundles.

Inputs follow SURVEY §8(d): joints ~ U[lower, upper] (what pinocchio.randomConfiguration
does for ``RobotState.randomize``, reference src/ramp/robot_state.py:142-154), free-flyer
position ~ U[-1, 1]^3 with a uniform unit quaternion, FP32, rows padded to a multiple of 4
floats.  Scenes: the reference capsule (reference examples/motion_planning.py:53-61) and the
box / sphere / capsule set of BASELINE config 2.
"""

from __future__ import annotations

import dataclasses

import numpy as np

from ramp_b200.kinematics import Model
from ramp_b200.robot_model import RobotModel, create_geometry_object, load_robot_model
from ramp_b200.robot_state import RobotState
from ramp_b200.se3 import SE3, Quaternion
from ramp_b200.shapes import Box, Capsule, Sphere


def reference_capsule():
    return create_geometry_object("capsule", Capsule(0.1, 0.4),
                                  SE3(Quaternion(0.707, 0.707, 0.0, 0.0), np.asarray([0.475, 0.0, 0.5])))


def obstacle_scene():
    return [
        create_geometry_object("box", Box(0.4, 0.4, 0.05), SE3(np.eye(3), np.asarray([0.5, 0.0, 0.2]))),
        create_geometry_object("sphere", Sphere(0.1), SE3(np.eye(3), np.asarray([0.4, 0.3, 0.5]))),
        reference_capsule(),
    ]


@dataclasses.dataclass
class Workload:
    name: str
    robot: str          # zoo configuration
    scene: str          # "none" | "capsule" | "obstacles"
    n: int              # configurations per step per GPU
    group: str | None
    note: str


WORKLOADS = {
    # BASELINE.json configs[0..4]; configs[2] (planning) is exercised by the edge workload
    "panda_4096": Workload("panda_4096", "panda/configs", "capsule", 4096, "arm",
                           "Panda stand-in, self + capsule scene, 4096 configs (latency regime)"),
    "ur5e_1m": Workload("ur5e_1m", "ur5e/configs", "obstacles", 1 << 20, "arm",
                        "UR5 stand-in, 1M configs, box/sphere/capsule scene"),
    "fr3_edges": Workload("fr3_edges", "fr3/configs", "capsule", 1 << 14, "arm",
                          "FR3 stand-in, planner edges through the batched motion validator"),
    "h1_16m": Workload("h1_16m", "unitree_h1/configs", "none", 1 << 24, None,
                       "H1 stand-in, floating base, self-collision, 16M configs"),
    "panda_64m": Workload("panda_64m", "panda/configs", "capsule", 1 << 26, "arm",
                          "Panda stand-in convex GJK, 64M configs sharded over the GPUs"),
    # the same robots with NON-capsule link hulls (rounded boxes, elbows, plates, blobs): the
    # inscribed capsules fill far less of such hulls, so many more pairs reach GJK
    "panda_boxy_64m": Workload("panda_boxy_64m", "panda_boxy/configs", "capsule", 1 << 26, "arm",
                               "Panda stand-in with boxy (non-capsule) link hulls, convex GJK, 64M configs"),
    "panda_boxy_4096": Workload("panda_boxy_4096", "panda_boxy/configs", "capsule", 4096, "arm",
                                "Panda stand-in with boxy link hulls, self + capsule scene, 4096 configs"),
    "h1_boxy_16m": Workload("h1_boxy_16m", "unitree_h1_boxy/configs", "none", 1 << 24, None,
                            "H1 stand-in with boxy link hulls, floating base, self-collision, 16M configs"),
    "rrr_float": Workload("rrr_float", "rrr/floating_configs", "capsule", 1 << 16, "arm",
                          "in-tree floating-base rrr (primitive shapes) with the reference capsule"),
}


def build(name: str) -> tuple[RobotModel, RobotState]:
    """Load the robot and add the scene objects the way the reference's examples do."""
    wl = WORKLOADS[name]
    rm = load_robot_model("zoo::" + wl.robot)
    rs = RobotState(rm)
    if wl.scene == "capsule":
        rs.add_object(reference_capsule())
    elif wl.scene == "obstacles":
        for obj in obstacle_scene():
            rs.add_object(obj)
    return rm, rs


def padded_nq(model: Model) -> int:
    return (model.nq + 3) & ~3


def sample_q(model: Model, n: int, seed: int, position_bound: float = 1.0) -> np.ndarray:
    """n random Pinocchio configurations as float32 [n, nq_pad] (vectorised restatement of
    kinematics.random_configuration)."""
    rng = np.random.default_rng(seed)
    q = np.zeros((n, padded_nq(model)), dtype=np.float32)
    lo, hi = model.lowerPositionLimit, model.upperPositionLimit

    def uni(i):
        a = lo[i] if np.isfinite(lo[i]) else -position_bound
        b = hi[i] if np.isfinite(hi[i]) else position_bound
        return rng.uniform(a, b, n)

    for j in model.joints[1:]:
        s, i = j.shortname(), j.idx_q
        if s.startswith("JointModelRUB") or s == "JointModelRevoluteUnboundedUnaligned":
            th = rng.uniform(-np.pi, np.pi, n)
            q[:, i], q[:, i + 1] = np.cos(th), np.sin(th)
        elif s == "JointModelPlanar":
            th = rng.uniform(-np.pi, np.pi, n)
            q[:, i], q[:, i + 1], q[:, i + 2], q[:, i + 3] = uni(i), uni(i + 1), np.cos(th), np.sin(th)
        elif s in ("JointModelFreeFlyer", "JointModelSpherical"):
            o = 3 if s == "JointModelFreeFlyer" else 0
            for k in range(o):
                q[:, i + k] = uni(i + k)
            v = rng.normal(size=(n, 4))
            q[:, i + o : i + o + 4] = v / np.linalg.norm(v, axis=1, keepdims=True)
        else:
            for k in range(j.nq):
                q[:, i + k] = uni(i + k)
    return q


def sample_q_torch(model: Model, n: int, seed: int, device, position_bound: float = 1.0):
    """Same distribution generated on the device (inputs of the large batches never cross
    PCIe)."""
    import torch

    g = torch.Generator(device=device)
    g.manual_seed(seed)
    q = torch.zeros((n, padded_nq(model)), dtype=torch.float32, device=device)
    lo, hi = model.lowerPositionLimit, model.upperPositionLimit

    def uni(i):
        a = lo[i] if np.isfinite(lo[i]) else -position_bound
        b = hi[i] if np.isfinite(hi[i]) else position_bound
        return torch.rand(n, generator=g, device=device) * float(b - a) + float(a)

    for j in model.joints[1:]:
        s, i = j.shortname(), j.idx_q
        if s.startswith("JointModelRUB") or s == "JointModelRevoluteUnboundedUnaligned":
            th = (torch.rand(n, generator=g, device=device) * 2 - 1) * np.pi
            q[:, i], q[:, i + 1] = torch.cos(th), torch.sin(th)
        elif s == "JointModelPlanar":
            th = (torch.rand(n, generator=g, device=device) * 2 - 1) * np.pi
            q[:, i], q[:, i + 1], q[:, i + 2], q[:, i + 3] = uni(i), uni(i + 1), torch.cos(th), torch.sin(th)
        elif s in ("JointModelFreeFlyer", "JointModelSpherical"):
            o = 3 if s == "JointModelFreeFlyer" else 0
            for k in range(o):
                q[:, i + k] = uni(i + k)
            v = torch.randn((n, 4), generator=g, device=device)
            q[:, i + o : i + o + 4] = v / v.norm(dim=1, keepdim=True)
        else:
            for k in range(j.nq):
                q[:, i + k] = uni(i + k)
    return q
