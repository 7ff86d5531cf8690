"""MotionPlanner with the reference's surface (reference src/ramp/motion_planner.py:247-625)
over the batched GPU validity path.

What replaces what:

* the per-state ``StateValidityCheckerFn`` callback chain (reference :469-483 ->
  robot_state.set_group_qpos -> is_state_valid -> check_collision) becomes ONE
  library call per edge: ``GpuMotionValidator.checkMotion`` interpolates at OMPL's default
  resolution, expands q, runs FK + narrow phase for every state of the edge and reduces the
  first invalid index on the device (ramp_check_edges);
* ``is_state_valid`` keeps its meaning (bounds and not in collision, reference :612-625);
* OMPL itself is a third-party wheel that is not installable here, so ``plan`` drives a
  small built-in RRTConnect (bidirectional, OMPL's default range of 20 % of the maximum
  extent, shortcut simplification, final interpolation) through the same validator
  interface an ``ompl.base.MotionValidator`` has.  When ``ompl`` is importable,
  ``make_ompl_motion_validator`` returns a real ``ob.MotionValidator`` subclass to install
  with ``si.setMotionValidator``.
"""

from __future__ import annotations

import logging
import time

import numpy as np
import torch

from ramp_b200 import spaces
from ramp_b200.ompl_codec import fill_ompl_state, from_ompl_state
from ramp_b200.robot_model import RobotModel
from ramp_b200.robot_state import RobotState

LOGGER = logging.getLogger(__name__)


class GpuMotionValidator:
    """``ompl::base::MotionValidator`` semantics on actuated-vector states.

    ``checkMotion(s1, s2)`` -> bool; ``checkMotion(s1, s2, last_valid)`` additionally fills
    ``last_valid`` (a 2-element list: state, t) like OMPL's pair overload.  Both follow from
    the kernel's output, the smallest invalid interpolation index j in [1, nd] (j = nd is s2
    itself): the bool is "no invalid j"; the last valid parameter is (j - 1) / nd."""

    def __init__(self, planner: "MotionPlanner"):
        self._p = planner
        self.valid_count = 0
        self.invalid_count = 0
        self.states_checked = 0

    def check_edges(self, q0, q1) -> dict:
        p = self._p
        p._sync_gpu()
        res = p._robot_model.gpu.check_edges(q0, q1, p.space)
        n_valid = int(res["valid"].sum().item())
        self.valid_count += n_valid
        self.invalid_count += int(res["valid"].numel()) - n_valid
        self.states_checked += int(res["nd"].clamp(min=1).sum().item())
        return res

    def checkMotion(self, s1, s2, last_valid: list | None = None) -> bool:  # noqa: N802
        p = self._p
        p._sync_gpu()
        res = p._robot_model.gpu.check_edges_host(np.asarray(s1, dtype=np.float32)[None], np.asarray(s2, dtype=np.float32)[None],
                                                  p.space)
        valid, first, nd = bool(res["valid"][0]), int(res["first_invalid"][0]), int(res["nd"][0])
        self.states_checked += max(nd, 1)
        if valid:
            self.valid_count += 1
            return True
        self.invalid_count += 1
        if last_valid is not None:
            t = (first - 1) / nd if nd > 0 else 0.0
            last_valid[:] = [spaces.interpolate(self._p.space, s1, s2, t), t]
        return False


class MotionPlanner:
    def __init__(self, robot_model: RobotModel, group_name: str) -> None:
        self._robot_model = robot_model
        self._group_name = group_name
        self.space = spaces.make_space(robot_model, group_name)
        self._reference_qpos = None
        self._qmap = None
        self.validator = GpuMotionValidator(self)
        self.rng = np.random.default_rng(0)

    # ------------------------------------------------------------------ state <-> values
    def as_state(self, robot_state: RobotState) -> np.ndarray:
        """Actuated values of the group (what ``as_ompl_state`` packs, reference :379-464)."""
        return np.asarray(robot_state.group_qpos(self._group_name), dtype=np.float64)

    def as_ompl_state(self, robot_state: RobotState, ompl_space, state=None):
        """The group's joint values as a compound OMPL state of ``ompl_space`` (reference :379-464).
        With the OMPL wheel present a new ``ob.State`` is built and returned; without it pass the
        state object to fill (anything indexable by sub-space like ``ob.State()``)."""
        values = robot_state.group_qpos(self._group_name)
        if state is not None:
            fill_ompl_state(ompl_space, state, values)
            return state
        from ompl import base as ob  # noqa: PLC0415 - optional dependency

        compound = ob.CompoundState(ompl_space)
        fill_ompl_state(ompl_space, compound(), values)
        return ob.State(compound)

    def state_validity_checker(self, reference_state: RobotState, ompl_space):
        """The ``(ob.State) -> bool`` callable the reference installs with
        ``setStateValidityChecker(ob.StateValidityCheckerFn(...))`` (reference :466-483): the state's
        joint values go into a private clone of ``reference_state``, then ``is_state_valid``."""
        reference_robot_state = reference_state.clone()

        def is_ompl_state_valid(state) -> bool:
            reference_robot_state.set_group_qpos(self._group_name, from_ompl_state(ompl_space, state))
            return self.is_state_valid(reference_robot_state)

        return is_ompl_state_valid

    def set_reference_state(self, reference_state: RobotState) -> None:
        """Every q slot outside the group comes from this state (the clone the reference
        captures in its validity closure, reference :466-483)."""
        self._reference_qpos = np.copy(reference_state.qpos)
        self._qmap = spaces.make_qmap(self._robot_model, self._group_name, self._reference_qpos)
        self._robot_model.gpu.set_qmap(self._qmap)

    def _sync_gpu(self) -> None:
        gpu = self._robot_model.gpu  # re-syncs the scene if objects were added
        if self._qmap is None:
            self.set_reference_state(RobotState(self._robot_model))
        elif gpu.qmap is not self._qmap:
            gpu.set_qmap(self._qmap)

    # ---------------------------------------------------------------------- validity
    def is_state_valid(self, robot_state: RobotState, *, verbose: bool = False) -> bool:
        """Within bounds and not in collision (reference :612-625)."""
        return spaces.satisfies_bounds(self.space, self.as_state(robot_state)) and not robot_state.check_collision(
            verbose=verbose)

    def are_states_valid(self, states) -> np.ndarray:
        """Batched ``is_state_valid`` on actuated vectors [N, n_act]."""
        self._sync_gpu()
        states = np.atleast_2d(np.asarray(states, dtype=np.float64))
        inb = spaces.satisfies_bounds_batch(self.space, states)
        hit = self._robot_model.gpu.check(states.astype(np.float32), actuated=True)["hit"].cpu().numpy()
        return inb & ~hit

    def project(self, states) -> np.ndarray:
        """TCP positions of a batch of actuated states: the projection
        ``ProjectionEvaluatorLinkPose`` evaluates one FK at a time (reference
        src/ramp/motion_planner.py:209-244,485-506), here one ramp_fk_batch call."""
        tcp = self._robot_model[self._group_name].tcp_link_name
        assert tcp is not None, f"tcp_link_name is not defined for group '{self._group_name}'"
        self._sync_gpu()
        model = self._robot_model.model
        frame = model.frames[model.getFrameId(tcp)]
        states = np.atleast_2d(np.asarray(states, dtype=np.float32))
        if frame.parentJoint == 0:
            return np.tile(frame.placement.translation, (len(states), 1))
        omi = self._robot_model.gpu.forward_kinematics(states, actuated=True)[:, frame.parentJoint - 1]
        p = torch.as_tensor(frame.placement.translation, dtype=torch.float32, device=omi.device)
        return (omi[:, :, :3] @ p + omi[:, :, 3]).cpu().numpy()

    # ---------------------------------------------------------------------- planning
    def _sample(self) -> np.ndarray:
        return spaces.sample_uniform(self.space, 1, self.rng, position_bound=10.0)[0]

    def _extend(self, tree: list, parents: list, target: np.ndarray, max_step: float):
        """One RRT extension toward ``target``; returns (status, new index): 0 trapped,
        1 advanced, 2 reached."""
        d = spaces.distance_many(self.space, tree, target)
        i = int(np.argmin(d))
        near, dist = tree[i], float(d[i])
        reached = dist <= max_step
        new = target if reached else spaces.interpolate(self.space, near, target, max_step / dist)
        if not self.validator.checkMotion(near, new):
            return 0, -1
        tree.append(np.asarray(new, dtype=np.float64))
        parents.append(i)
        return (2 if reached else 1), len(tree) - 1

    def _rrt_connect(self, start: np.ndarray, goal: np.ndarray, timeout: float):
        max_step = 0.2 * spaces.maximum_extent(self.space)  # OMPL's default planner range
        ta, pa, tb, pb = [start], [-1], [goal], [-1]
        a_is_start = True
        t_end = time.perf_counter() + timeout
        # a straight connection first (what RRTConnect finds in its first iterations)
        while time.perf_counter() < t_end:
            rand = self._sample()
            st, ia = self._extend(ta, pa, rand, max_step)
            if st != 0:
                target = ta[ia]
                while True:  # connect: keep extending the other tree toward the new node
                    st2, ib = self._extend(tb, pb, target, max_step)
                    if st2 != 1:
                        break
                if st2 == 2:
                    def chain(tree, par, i):
                        out = []
                        while i >= 0:
                            out.append(tree[i])
                            i = par[i]
                        return out
                    left, right = chain(ta, pa, ia)[::-1], chain(tb, pb, ib)[1:]
                    path = left + right
                    return path if a_is_start else path[::-1]
            ta, pa, tb, pb = tb, pb, ta, pa
            a_is_start = not a_is_start
        return None

    def _shortcut(self, path: list, rounds: int = 30) -> list:
        """Random shortcutting (the role of simplifySolution / ropeShortcutPath, reference
        :570-584): replace a sub-path by a straight edge when that edge is valid."""
        path = list(path)
        for _ in range(rounds):
            if len(path) <= 2:
                break
            i, j = sorted(self.rng.choice(len(path), size=2, replace=False))
            if j - i >= 2 and self.validator.checkMotion(path[i], path[j]):
                path = path[: i + 1] + path[j:]
        return path

    def _interpolate_path(self, path: list) -> list:
        """PathGeometric::interpolate: one state per valid segment of every edge."""
        out = [path[0]]
        for a, b in zip(path[:-1], path[1:]):
            nd = max(spaces.valid_segment_count(self.space, a, b), 1)
            out.extend(spaces.interpolate(self.space, a, b, j / nd) for j in range(1, nd + 1))
        return out

    def check_path(self, path: list) -> bool:
        """PathGeometric::check in one batched call over all edges of the path."""
        if len(path) < 2:
            return True
        res = self.validator.check_edges(np.asarray(path[:-1], dtype=np.float32), np.asarray(path[1:], dtype=np.float32))
        return bool(res["valid"].all().item())

    def path_clearance(self, path: list) -> np.ndarray:
        """Clearance (smallest signed distance over the enabled pairs) at every interpolated state
        of the path, one batched call -- the state cost of the reference's PathClearanceObjective
        (src/ramp/motion_planner.py:165-206, dead code there because it is never installed)."""
        self._sync_gpu()
        states = np.asarray(self._interpolate_path(path), dtype=np.float32)
        return self._robot_model.gpu.clearance(states, actuated=True).cpu().numpy()

    def plan(self, start_state: RobotState, group_goal_qpos, timeout: float = 1.0,
             planner: str | None = None) -> list[RobotState] | None:
        """reference :509-610.  ``planner`` accepts None / "RRTConnect" (the built-in)."""
        idx = self._robot_model[self._group_name].joint_position_indices
        assert len(group_goal_qpos) == len(idx)
        if planner not in (None, "RRTConnect"):
            raise AttributeError(f"Planner '{planner}' not found - Available planners: ['RRTConnect']")
        goal_state = start_state.clone()
        goal_state.set_group_qpos(self._group_name, group_goal_qpos)
        self.set_reference_state(start_state)
        if not self.is_state_valid(start_state, verbose=True):
            LOGGER.error("Start state is invalid - in collision or out of bounds")
            return None
        if not self.is_state_valid(goal_state, verbose=True):
            LOGGER.error("Goal state is invalid - in collision or out of bounds")
            return None
        path = self._rrt_connect(self.as_state(start_state), self.as_state(goal_state), timeout)
        if path is None:
            LOGGER.info("Did not find solution!")
            return None
        if not self.check_path(path):
            LOGGER.warning("Path fails check!")
        path = self._shortcut(path)
        path = self._interpolate_path(path)
        if not self.check_path(path):
            LOGGER.warning("Interpolated simplified path fails check!")
        solution = []
        ref = start_state.clone()
        for s in path:
            ref.set_group_qpos(self._group_name, s)
            solution.append(ref.clone())
        LOGGER.info(f"Found solution with {len(solution)} waypoints")
        return solution


def make_ompl_motion_validator(planner: MotionPlanner, si, ob=None):
    """An ``ompl.base.MotionValidator`` over the GPU validator, for installs that have the
    OMPL wheel: ``si.setMotionValidator(make_ompl_motion_validator(planner, si))`` right after
    ``og.SimpleSetup`` is built (reference motion_planner.py:366).  States are decoded with
    ``ompl_codec.from_ompl_state``; ``ob`` is the ``ompl.base`` module (imported when omitted; the
    tests pass a stand-in with a ``MotionValidator`` base class)."""
    if ob is None:
        from ompl import base as ob  # noqa: PLC0415 - optional dependency

    class _Validator(ob.MotionValidator):
        def __init__(self):
            super().__init__(si)

        def checkMotion(self, s1, s2, last_valid=None):  # noqa: N802
            space = si.getStateSpace()
            a, b = from_ompl_state(space, s1), from_ompl_state(space, s2)
            if last_valid is None:
                return planner.validator.checkMotion(a, b)
            lv: list = []
            ok = planner.validator.checkMotion(a, b, lv)
            if not ok:
                last_valid.second = lv[1]
                if last_valid.first is not None:
                    space.interpolate(s1, s2, lv[1], last_valid.first)
            return ok

    return _Validator()
