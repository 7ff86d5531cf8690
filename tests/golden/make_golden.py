"""Generates the fixtures in this directory.  Run in the authoring container, where the
reference tree is mounted at /root/reference (it is not on the GPU box):

    python tests/golden/make_golden.py

* ``reference_tables.json`` -- the kinematic / geometry / pair tables obtained by loading the
  REFERENCE's own robot files (robots/**/configs.toml, tests/double_pendulum.toml,
  robots/*.urdf) with this repo's description readers.  tests/test_golden.py checks that the
  self-contained zoo (ramp_b200/zoo.py) gives the same tables anywhere, without the reference.
* ``oracle_vectors.npz`` -- seeded configurations and what the FP64 oracle says about them
  (flags, first pair, per-pair signed distance, joint poses): a regression anchor for the oracle
  and known answers for the GPU path.  The reference's own arithmetic (Pinocchio / coal) cannot
  run here, so these are oracle outputs, not reference outputs.
"""

import json
import sys
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
sys.path.insert(0, str(HERE.parents[1]))
REF = Path("/root/reference")

REF_FILES = [("rrr/configs", "robots/rrr/configs.toml"), ("rrr/planar_configs", "robots/rrr/planar_configs.toml"),
             ("rrr/floating_configs", "robots/rrr/floating_configs.toml"), ("planar_rrr/configs", "robots/planar_rrr/configs.toml"),
             ("rrr_mj/configs", "robots/rrr_mj/configs.toml"), ("acrobot/configs", "robots/acrobot/configs.toml"),
             ("double_pendulum", "tests/double_pendulum.toml"), ("floating", "robots/floating.urdf"), ("planar", "robots/planar.urdf")]
VECTORS = [("rrr/configs", "capsule"), ("rrr/floating_configs", "capsule"), ("panda/configs", "capsule"), ("ur5e/configs", "obstacles"),
           ("panda_boxy/configs", "capsule"), ("unitree_h1/configs", "none"), ("joint_zoo", "capsule")]
IK_VECTORS = ["panda/configs", "ur5e/configs"]


def tables(rm) -> dict:
    from ramp_b200.desc import ModelDesc

    d = ModelDesc(rm.model, rm.collision_model)
    return {
        "joints": [[int(j.type), int(j.parent), int(j.idx_q), [float(x) for x in j.axis], [round(float(x), 12) for x in j.placement]]
                   for j in d.joints[: d.struct.n_joints]],
        "geoms": [[int(g.type), int(g.parent_joint), [float(x) for x in g.params], [round(float(x), 12) for x in g.placement]]
                  for g in d.geoms[: d.n_geoms]],
        "pairs": np.asarray(d.pairs).tolist(),
        "geometry_names": [g.name for g in rm.collision_model.geometryObjects],
        "frames": [[f.name, int(f.parentJoint), int(f.type)] for f in rm.model.frames],
        "joint_names": list(rm.model.names),
        "groups": {g: {"joints": list(rm[g].joints), "tcp": rm[g].tcp_link_name,
                       "named_states": {k: np.asarray(v).tolist() for k, v in rm[g].named_states.items()}} for g in rm.groups},
    }


def scene_objects(name):
    from ramp_b200 import workloads

    return {"capsule": [workloads.reference_capsule()], "obstacles": workloads.obstacle_scene(), "none": []}[name]


def main():
    from oracle import oracle
    from ramp_b200 import workloads
    from ramp_b200.desc import ModelDesc
    from ramp_b200.robot_model import load_robot_model
    from ramp_b200.robot_state import RobotState

    ref = {zoo: tables(load_robot_model(REF / path)) for zoo, path in REF_FILES}
    (HERE / "reference_tables.json").write_text(json.dumps(ref, indent=0, sort_keys=True))
    out = {}
    for zoo, scene in VECTORS:
        rm = load_robot_model("zoo::" + zoo)
        rs = RobotState(rm)
        for o in scene_objects(scene):
            rs.add_object(o)
        q = workloads.sample_q(rm.model, 96, seed=1234)[:, : rm.model.nq].astype(np.float64)
        d = ModelDesc(rm.model, rm.collision_model)
        res = oracle.check(d, q)
        key = zoo.replace("/", "__")
        out[key + "__q"] = q
        out[key + "__hit"] = res["hit"].astype(np.uint8)
        out[key + "__first_pair"] = res["first_pair"].astype(np.int32)
        out[key + "__pair_dist"] = res["pair_dist"]
        out[key + "__fk"] = oracle.fk(d, q[:8])
    # batched IK: seeds, targets and what the FP64 iteration returns after 5 steps and at convergence
    from ramp_b200 import spaces
    from ramp_b200.robot_state import RobotState as RS

    for zoo in IK_VECTORS:
        rm = load_robot_model("zoo::" + zoo)
        d = ModelDesc(rm.model, rm.collision_model)
        qmap = spaces.make_qmap(rm, "arm", RS.from_named_state(rm, "arm", "home").qpos)
        params = rm.ik_params("arm")
        rng = np.random.default_rng(77)
        lo, hi = np.maximum(params.lower, -3.0), np.minimum(params.upper, 3.0)
        goal = rng.uniform(lo, hi, (64, qmap.n_act))
        frame = rm.model.frames[rm.model.getFrameId(rm["arm"].tcp_link_name)]
        omi = oracle.fk(d, oracle.qmap_apply(qmap, goal))[:, frame.parentJoint - 1]
        pl = frame.placement.rt34()
        targets = np.concatenate([omi[:, :, :3] @ pl[:, :3], (omi[:, :, :3] @ pl[:, 3] + omi[:, :, 3])[:, :, None]], axis=2)
        seeds = np.clip(goal + rng.normal(scale=0.3, size=goal.shape), params.lower, params.upper).astype(np.float32).astype(np.float64)
        five = oracle.ik(d, qmap, rm.ik_params("arm", max_iterations=5, tolerance=1e-12), seeds, targets)
        full = oracle.ik(d, qmap, params, seeds, targets)
        key = "ik__" + zoo.replace("/", "__")
        out[key + "__seeds"], out[key + "__targets"] = seeds, targets
        out[key + "__err5"], out[key + "__ok"], out[key + "__q"] = five["err"], full["ok"].astype(np.uint8), full["q"]
    np.savez_compressed(HERE / "oracle_vectors.npz", **out)
    print("wrote", HERE / "reference_tables.json", (HERE / "reference_tables.json").stat().st_size, "B;",
          HERE / "oracle_vectors.npz", (HERE / "oracle_vectors.npz").stat().st_size, "B")


if __name__ == "__main__":
    main()
