"""RobotState with the reference's surface (reference src/ramp/robot_state.py:51-392); the
queries that were Pinocchio + coal calls run on the GPU through ``robot_model.gpu``.

Single-state calls use the batch path with N = 1.  ``differential_ik`` /
``jacobian`` are out of scope (QP-based IK is not a batch path).
"""

from __future__ import annotations

import logging

import numpy as np

from ramp_b200.exceptions import FrameNotFound
from ramp_b200.kinematics import CollisionPair, GeometryObject, neutral, random_configuration
from ramp_b200.robot_model import (
    RobotModel,
    as_pinocchio_qpos,
    get_converted_qpos,
    set_positions,
)
from ramp_b200.se3 import SE3

LOGGER = logging.getLogger(__name__)


def print_collision_results(collision_model, pair_mask) -> None:
    """reference src/ramp/robot_state.py:31-48."""
    for k, cp in enumerate(collision_model.collisionPairs):
        if pair_mask[k]:
            LOGGER.debug(
                f"Collision between: ({collision_model.geometryObjects[cp.first].name}"
                f", {collision_model.geometryObjects[cp.second].name})")


class RobotState:
    def __init__(self, robot_model: RobotModel, qpos: np.ndarray | None = None, qvel: np.ndarray | None = None):
        assert qpos is None or len(qpos) == robot_model.model.nq
        self.robot_model = robot_model
        self.qpos = neutral(robot_model.model) if qpos is None else np.asarray(qpos, dtype=np.float64)
        self.qvel = np.zeros(robot_model.model.nv) if qvel is None else qvel
        self.geometry_objects: dict[str, int] = {}

    def clone(self) -> "RobotState":
        rs = RobotState(self.robot_model, np.copy(self.qpos), np.copy(self.qvel))
        rs.geometry_objects = self.geometry_objects.copy()
        return rs

    # ------------------------------------------------------------- q bookkeeping
    def joint_qpos(self, joint_name: str) -> np.ndarray:
        from ramp_b200.robot_model import joint_indices

        model = self.robot_model.model
        qpos = get_converted_qpos(self.robot_model, self.qpos)
        return qpos[joint_indices(model.joints[model.getJointId(joint_name)])]

    def set_joint_qpos(self, joint_name: str, value) -> None:
        from ramp_b200.robot_model import joint_indices

        model = self.robot_model.model
        indices = joint_indices(model.joints[model.getJointId(joint_name)])
        assert len(value) == len(indices), f"Expected {len(indices)} joint positions, got {len(value)}"
        set_positions(self.robot_model, self.qpos, indices, value)

    def group_qpos(self, group_name: str) -> np.ndarray:
        return get_converted_qpos(self.robot_model, self.qpos)[self.robot_model[group_name].joint_position_indices]

    def set_group_qpos(self, group_name: str, value) -> None:
        idx = self.robot_model[group_name].joint_position_indices
        assert len(value) == len(idx), f"Expected {len(idx)} joint positions, got {len(value)}"
        set_positions(self.robot_model, self.qpos, idx, value)

    def actuated_qpos(self) -> np.ndarray:
        qpos = get_converted_qpos(self.robot_model, self.qpos)
        return np.concatenate([qpos[g.joint_position_indices] for g in self.robot_model.groups.values()])

    def set_actuated_qpos(self, value) -> None:
        idx = np.concatenate([g.joint_position_indices for g in self.robot_model.groups.values()])
        set_positions(self.robot_model, self.qpos, idx, np.asarray(value, dtype=np.float64)[: len(idx)])

    def randomize(self, group_name: str | None = None, rng: np.random.Generator | None = None) -> None:
        qpos = get_converted_qpos(self.robot_model, random_configuration(self.robot_model.model, rng))
        groups = [(group_name, self.robot_model[group_name])] if group_name else self.robot_model.groups.items()
        for name, group in groups:
            self.set_group_qpos(name, qpos[group.joint_position_indices])

    @classmethod
    def from_random(cls, robot_model: RobotModel, rng: np.random.Generator | None = None):
        rs = cls(robot_model)
        rs.randomize(rng=rng)
        return rs

    @classmethod
    def from_actuated_qpos(cls, robot_model: RobotModel, joint_positions):
        rs = cls(robot_model)
        rs.set_actuated_qpos(joint_positions)
        return rs

    @classmethod
    def from_pinocchio_qpos(cls, robot_model: RobotModel, q):
        assert len(q) == robot_model.model.nq
        return cls(robot_model, q)

    @classmethod
    def from_named_state(cls, robot_model: RobotModel, group_name: str, state_name: str):
        return cls(robot_model, as_pinocchio_qpos(robot_model, neutral(robot_model.model), group_name,
                                                  robot_model[group_name].named_states[state_name]))

    # ------------------------------------------------------------------ GPU queries
    def get_frame_pose(self, target_frame_name: str) -> SE3:
        """oMf = oMi[parentJoint] * frame.placement (reference robot_state.py:194-212)."""
        model = self.robot_model.model
        fid = model.getFrameId(target_frame_name)
        if fid >= len(model.frames):
            raise FrameNotFound(target_frame_name, model.frames)
        frame = model.frames[fid]
        if frame.parentJoint == 0:
            return frame.placement.copy()
        omi = self.robot_model.gpu.forward_kinematics(self.qpos[None].astype(np.float32))[0].cpu().numpy()
        return SE3.from_rt34(omi[frame.parentJoint - 1].astype(np.float64)) * frame.placement

    def differential_ik(self, group_name: str, target_pose, iteration_callback=None) -> bool:
        """Move the group so that its tool frame reaches ``target_pose`` ([x, y, z, qx, qy, qz, qw] or
        a 4x4 matrix); True and the state updated on success, False and the state untouched
        otherwise (reference src/ramp/robot_state.py:237-304).  One seed -- this state -- through
        the batched damped-least-squares kernel (ramp_ik_batch); ``iteration_callback`` is accepted
        for signature compatibility and called once with the final state (the iterations run on
        the device)."""
        pose = np.asarray(target_pose, dtype=np.float64)
        if pose.shape == (4, 4):
            target = pose[:3, :]
        else:
            assert pose.shape == (7,), "target pose: [x, y, z, qx, qy, qz, qw] or a 4x4 matrix"
            from ramp_b200.se3 import Quaternion

            qn = pose[3:] / np.linalg.norm(pose[3:])
            target = np.concatenate([Quaternion(qn[3], qn[0], qn[1], qn[2]).matrix(), pose[:3, None]], axis=1)
        res = self.robot_model.differential_ik_batch(group_name, self.group_qpos(group_name)[None].astype(np.float32), target,
                                                     reference_state=self)
        ok = bool(res["ok"][0].item())
        if ok:
            self.set_group_qpos(group_name, res["q"][0].cpu().numpy().astype(np.float64))
        if iteration_callback is not None:
            iteration_callback(self.clone())
        return ok

    def add_object(self, geometry_object: GeometryObject) -> None:
        """Append a universe-attached object and pair it with every ROBOT geometry (objects
        added earlier are excluded, so scene-scene pairs are never checked) -- reference
        robot_state.py:307-331, including its mutation of the shared RobotModel."""
        cm = self.robot_model.collision_model
        gid = cm.addGeometryObject(geometry_object)
        self.geometry_objects[geometry_object.name] = gid
        self.robot_model.visual_model.addGeometryObject(geometry_object)
        for geometry_id in range(len(cm.geometryObjects) - len(self.geometry_objects)):
            cm.addCollisionPair(CollisionPair(gid, geometry_id))

    def compute_distance(self, object1: str, object2: str) -> float:
        """Signed distance of one pair (reference robot_state.py:333-364)."""
        cm = self.robot_model.collision_model
        pair_idx = cm.findCollisionPair(CollisionPair(cm.getGeometryId(object1), cm.getGeometryId(object2)))
        if pair_idx >= len(cm.collisionPairs):
            raise ValueError(f"({object1}, {object2}) is not a collision pair")
        return float(self.robot_model.gpu.distance(self.qpos[None].astype(np.float32), [pair_idx])[0, 0])

    def check_collision(self, *, verbose: bool = False) -> bool:
        """True if any enabled pair collides (reference robot_state.py:366-392).  ``verbose``
        evaluates every pair (no early-out) and logs the colliding ones at DEBUG."""
        res = self.robot_model.gpu.check_host(self.qpos.astype(np.float32), pair_mask=verbose)
        if verbose:
            print_collision_results(self.robot_model.collision_model, res["pair_mask"][0])
        return bool(res["hit"][0])
