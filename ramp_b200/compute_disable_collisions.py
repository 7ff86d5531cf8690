"""Disable-collision helpers with the reference's surface
(reference src/ramp/compute_disable_collisions.py:16-226); ``get_collision_pairs`` runs the
"all pairs" mode of the GPU kernel (computeCollisions with stop_at_first_collision=False)."""

from __future__ import annotations

import enum
import itertools
import logging

import numpy as np

from ramp_b200.kinematics import BODY, SIZE_T_MAX, CollisionPair, neutral, remove_collision_pairs_from_srdf

LOGGER = logging.getLogger(__name__)


class DisabledReason(enum.StrEnum):
    DEFAULT = "Default"
    ADJACENT = "Adjacent"
    ALWAYS = "Always"
    NEVER = "Never"


def get_collision_pair_names(model, collision_model, pair) -> tuple[str, str]:
    """Body-frame names of a pair, geometry names for frame-less objects (reference :38-53)."""
    out = []
    for gid in (pair.first, pair.second):
        obj = collision_model.geometryObjects[gid]
        out.append(obj.name if obj.parentFrame == SIZE_T_MAX else model.frames[obj.parentFrame].name)
    return out[0], out[1]


def disable_collision_pair(robot_model, link1: str, link2: str) -> None:
    """Remove every pair between the geometries of two links / named objects (reference :56-89)."""
    cm = robot_model.collision_model
    g1 = [g.name for g in robot_model.body_geometries(link1)] or [link1]
    g2 = [g.name for g in robot_model.body_geometries(link2)] or [link2]
    for a, b in itertools.product(g1, g2):
        cm.removeCollisionPair(CollisionPair(cm.getGeometryId(a), cm.getGeometryId(b)))


def disable_collision(robot_model, disable_collisions, *, verbose=False) -> None:
    """Remove the pairs of a [(link1, link2, reason)] list, SRDF semantics (reference :92-136)."""
    if verbose and disable_collisions:
        LOGGER.info("Disabling: %s", disable_collisions)
    remove_collision_pairs_from_srdf(robot_model.model, robot_model.collision_model,
                                     [(a, b) for a, b, _ in disable_collisions])


def get_collision_pairs(robot_model, qpos) -> dict:
    """{(name1, name2): CollisionPair} of every pair colliding at ``qpos`` (reference :139-174)."""
    res = robot_model.gpu.check_host(np.asarray(qpos, dtype=np.float32), pair_mask=True)
    cm = robot_model.collision_model
    return {get_collision_pair_names(robot_model.model, cm, cm.collisionPairs[k]): cm.collisionPairs[k]
            for k in np.flatnonzero(res["pair_mask"][0])}


def default_collisions(robot_model) -> list:
    """Pairs colliding at the neutral configuration (reference :178-196)."""
    pairs = get_collision_pairs(robot_model, neutral(robot_model.model))
    return [(a, b, DisabledReason.DEFAULT) for a, b in pairs]


def adjacent_collisions(robot_model) -> list:
    """(parent body, body, ADJACENT) for every geometry-bearing link (reference :200-226)."""
    model = robot_model.model

    def body_parent(frame):
        while (parent := model.frames[frame.parentFrame]).type != BODY and frame.name != "universe":
            frame = parent
        return parent

    out = set()
    for obj in robot_model.collision_model.geometryObjects:
        if obj.parentFrame == SIZE_T_MAX:
            continue
        cur = model.frames[obj.parentFrame]
        par = body_parent(cur)
        if par.name == "universe" or cur.name == "universe":
            continue
        out.add((par.name, cur.name, DisabledReason.ADJACENT))
    return list(out)


def sampled_collisions(robot_model, n_samples: int = 1 << 17, seed: int = 0, batch: int = 1 << 17) -> list:
    """ALWAYS / NEVER in collision, the two reasons the reference leaves as TODOs
    (reference src/ramp/compute_disable_collisions.py:229-230): the all-pairs mode of the
    validity kernel over ``n_samples`` random configurations, reduced per pair.  Returns
    [(link1, link2, DisabledReason.ALWAYS | NEVER), ...] for the pairs currently enabled."""
    import torch

    from ramp_b200 import workloads

    gpu = robot_model.gpu
    cm = robot_model.collision_model
    npairs = len(cm.collisionPairs)
    if npairs == 0:
        return []
    any_hit = torch.zeros(npairs, dtype=torch.bool, device=gpu.device)
    all_hit = torch.ones(npairs, dtype=torch.bool, device=gpu.device)
    done = 0
    while done < n_samples:
        n = min(batch, n_samples - done)
        q = workloads.sample_q_torch(robot_model.model, n, seed=seed + done, device=gpu.device)
        mask = gpu.check(q, pair_mask=True)["pair_mask"]
        any_hit |= mask.any(dim=0)
        all_hit &= mask.all(dim=0)
        done += n
    any_hit, all_hit = any_hit.cpu().numpy(), all_hit.cpu().numpy()
    out = []
    for k, pair in enumerate(cm.collisionPairs):
        names = get_collision_pair_names(robot_model.model, cm, pair)
        if all_hit[k]:
            out.append((*names, DisabledReason.ALWAYS))
        elif not any_hit[k]:
            out.append((*names, DisabledReason.NEVER))
    return out
