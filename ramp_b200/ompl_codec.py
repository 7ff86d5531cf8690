"""OMPL compound states <-> actuated joint values (reference src/ramp/motion_planner.py:89-162
``from_ompl_state`` and :379-464 ``as_ompl_state``).

Both directions only index the state objects (``state[i]``, ``.value``, ``.getX()``,
``.rotation().x`` ...), so they are written against that duck-typed surface: with the OMPL wheel
present they take ``ob.CompoundStateSpace`` / ``ob.State`` objects, without it (this image) the
tests drive them with stand-in classes of the same shape.  Sub-space kinds are told apart by
``getType()``; the numbers are OMPL's ``StateSpaceTypes.h`` enumeration and are re-read from the
``ompl`` module when it is importable.

Covered: RealVector, SO2, SO3, SE2 (and the Dubins / Reeds-Shepp spaces, whose states are SE2
states), SE3 -- the spaces reference motion_planner.py:264-359 builds for revolute, prismatic,
continuous, planar and floating joints.  The Owen / Vana spaces (fixed-wing path families) are
outside this build's scope and raise ``ValueError("Unsupported space")`` like any unknown kind.
"""

from __future__ import annotations

STATE_SPACE_REAL_VECTOR, STATE_SPACE_SO2, STATE_SPACE_SO3, STATE_SPACE_SE2, STATE_SPACE_SE3 = 1, 2, 3, 4, 5
STATE_SPACE_DUBINS, STATE_SPACE_REEDS_SHEPP = 8, 9


def _kinds() -> dict[str, int]:
    names = ("STATE_SPACE_REAL_VECTOR", "STATE_SPACE_SO2", "STATE_SPACE_SO3", "STATE_SPACE_SE2", "STATE_SPACE_SE3",
             "STATE_SPACE_DUBINS", "STATE_SPACE_REEDS_SHEPP")
    out = {n: globals()[n] for n in names}
    try:
        from ompl import base as ob  # noqa: PLC0415 - optional dependency
    except ImportError:
        return out
    for n in names:
        if hasattr(ob, n):
            out[n] = int(getattr(ob, n))
    return out


def _is(kind: int, k: dict, *names: str) -> bool:
    return any(int(kind) == k[n] for n in names)


def from_ompl_state(space, state) -> list[float]:
    """Compound OMPL state -> actuated joint values, sub-space by sub-space
    (reference src/ramp/motion_planner.py:89-162)."""
    k = _kinds()
    joint_positions: list[float] = []
    for space_idx in range(space.getSubspaceCount()):
        subspace = space.getSubspace(space_idx)
        substate = state[space_idx]
        kind = subspace.getType()
        if _is(kind, k, "STATE_SPACE_REAL_VECTOR"):
            joint_positions.extend(float(substate[i]) for i in range(subspace.getDimension()))
        elif _is(kind, k, "STATE_SPACE_SO3"):
            joint_positions.extend([substate.x, substate.y, substate.z, substate.w])
        elif _is(kind, k, "STATE_SPACE_SE2", "STATE_SPACE_DUBINS", "STATE_SPACE_REEDS_SHEPP"):
            joint_positions.extend([substate.getX(), substate.getY(), substate.getYaw()])
        elif _is(kind, k, "STATE_SPACE_SO2"):
            joint_positions.append(substate.value)
        elif _is(kind, k, "STATE_SPACE_SE3"):
            rotation = substate.rotation()
            joint_positions.extend([substate.getX(), substate.getY(), substate.getZ(),
                                    rotation.x, rotation.y, rotation.z, rotation.w])
        else:
            msg = f"Unsupported space: {subspace}"
            raise ValueError(msg)
    return [float(v) for v in joint_positions]


def fill_ompl_state(space, internal_state, joint_positions) -> None:
    """Actuated joint values -> the sub-states of a compound OMPL state, in place
    (reference src/ramp/motion_planner.py:379-464).  ``internal_state`` is what ``ob.State()``
    returns (indexable by sub-space)."""
    k = _kinds()
    i = 0
    for space_idx in range(space.getSubspaceCount()):
        subspace = space.getSubspace(space_idx)
        substate = internal_state[space_idx]
        kind = subspace.getType()
        if _is(kind, k, "STATE_SPACE_SO2"):
            substate.value = float(joint_positions[i])
            i += 1
        elif _is(kind, k, "STATE_SPACE_SE2", "STATE_SPACE_DUBINS", "STATE_SPACE_REEDS_SHEPP"):
            substate.setX(float(joint_positions[i]))
            substate.setY(float(joint_positions[i + 1]))
            substate.setYaw(float(joint_positions[i + 2]))
            i += 3
        elif _is(kind, k, "STATE_SPACE_SO3"):
            substate.x, substate.y, substate.z, substate.w = (float(v) for v in joint_positions[i:i + 4])
            i += 4
        elif _is(kind, k, "STATE_SPACE_REAL_VECTOR"):
            for idx in range(subspace.getDimension()):
                substate[idx] = float(joint_positions[i + idx])
            i += subspace.getDimension()
        elif _is(kind, k, "STATE_SPACE_SE3"):
            substate.setX(float(joint_positions[i]))
            substate.setY(float(joint_positions[i + 1]))
            substate.setZ(float(joint_positions[i + 2]))
            rotation = substate.rotation()
            rotation.x, rotation.y, rotation.z, rotation.w = (float(v) for v in joint_positions[i + 3:i + 7])
            i += 7
        else:
            msg = f"Unsupported space: {subspace}"
            raise ValueError(msg)
    assert i == len(joint_positions), f"{len(joint_positions)} joint values for a space of {i}"
