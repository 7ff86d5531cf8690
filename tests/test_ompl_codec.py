"""from_ompl_state / as_ompl_state / the state-validity callable against stand-in state classes
shaped like OMPL's (reference src/ramp/motion_planner.py:89-162, :379-464, :466-483).  OMPL is a
third-party wheel that is not installable here; the codecs only index state objects, so classes
with the same accessors exercise every branch."""

import numpy as np
import pytest

from ramp_b200 import ompl_codec as oc


class _Sub:
    def __init__(self, kind, dim=0):
        self.kind, self.dim = kind, dim

    def getType(self):  # noqa: N802
        return self.kind

    def getDimension(self):  # noqa: N802
        return self.dim

    def __repr__(self):
        return f"<space kind {self.kind}>"


class _Compound:
    def __init__(self, subs):
        self.subs = subs

    def getSubspaceCount(self):  # noqa: N802
        return len(self.subs)

    def getSubspace(self, i):  # noqa: N802
        return self.subs[i]

    def getSubspaces(self):  # noqa: N802
        return list(self.subs)


class _Real(list):
    pass


class _SO2:
    value = 0.0


class _Quat:
    x = y = z = 0.0
    w = 1.0


class _SE2:
    def __init__(self):
        self.v = [0.0, 0.0, 0.0]

    def getX(self): return self.v[0]  # noqa: N802,E704
    def getY(self): return self.v[1]  # noqa: N802,E704
    def getYaw(self): return self.v[2]  # noqa: N802,E704
    def setX(self, x): self.v[0] = x  # noqa: N802,E704
    def setY(self, y): self.v[1] = y  # noqa: N802,E704
    def setYaw(self, a): self.v[2] = a  # noqa: N802,E704


class _SE3:
    def __init__(self):
        self.t, self.q = [0.0, 0.0, 0.0], _Quat()

    def getX(self): return self.t[0]  # noqa: N802,E704
    def getY(self): return self.t[1]  # noqa: N802,E704
    def getZ(self): return self.t[2]  # noqa: N802,E704
    def setX(self, x): self.t[0] = x  # noqa: N802,E704
    def setY(self, y): self.t[1] = y  # noqa: N802,E704
    def setZ(self, z): self.t[2] = z  # noqa: N802,E704
    def rotation(self): return self.q  # noqa: E704


def make_state(space):
    mk = {oc.STATE_SPACE_REAL_VECTOR: lambda s: _Real([0.0] * s.dim), oc.STATE_SPACE_SO2: lambda s: _SO2(),
          oc.STATE_SPACE_SO3: lambda s: _Quat(), oc.STATE_SPACE_SE2: lambda s: _SE2(), oc.STATE_SPACE_SE3: lambda s: _SE3(),
          oc.STATE_SPACE_DUBINS: lambda s: _SE2(), oc.STATE_SPACE_REEDS_SHEPP: lambda s: _SE2()}
    return [mk[s.kind](s) for s in space.subs]


def space_for(planner_space):
    """The compound OMPL space reference motion_planner.py:264-359 would build for our descriptor."""
    kind = {0: oc.STATE_SPACE_REAL_VECTOR, 1: oc.STATE_SPACE_SO2, 2: oc.STATE_SPACE_SE2, 3: oc.STATE_SPACE_SE3, 4: oc.STATE_SPACE_SO3}
    return _Compound([_Sub(kind[s["type"]], s.get("dim", 0)) for s in planner_space.subspaces])


def test_every_space_kind_round_trips():
    space = _Compound([_Sub(oc.STATE_SPACE_REAL_VECTOR, 1), _Sub(oc.STATE_SPACE_SE3), _Sub(oc.STATE_SPACE_SO2),
                       _Sub(oc.STATE_SPACE_SE2), _Sub(oc.STATE_SPACE_REAL_VECTOR, 3), _Sub(oc.STATE_SPACE_SO3),
                       _Sub(oc.STATE_SPACE_DUBINS), _Sub(oc.STATE_SPACE_REEDS_SHEPP)])
    values = list(np.round(np.random.default_rng(0).uniform(-1, 1, 1 + 7 + 1 + 3 + 3 + 4 + 3 + 3), 6))
    state = make_state(space)
    oc.fill_ompl_state(space, state, values)
    assert state[1].rotation().w == values[7] and state[2].value == values[8] and state[3].getYaw() == values[11]
    assert oc.from_ompl_state(space, state) == values
    with pytest.raises(AssertionError):
        oc.fill_ompl_state(space, make_state(space), values + [0.0])
    bad = _Compound([_Sub(77)])
    with pytest.raises(ValueError, match="Unsupported space"):
        oc.from_ompl_state(bad, [None])
    with pytest.raises(ValueError, match="Unsupported space"):
        oc.fill_ompl_state(bad, [None], [0.0])


@pytest.mark.parametrize("cfg", ["rrr/configs", "rrr/planar_configs", "rrr/floating_configs", "planar_rrr/configs"])
def test_group_values_round_trip_through_a_planner(cfg):
    """MotionPlanner.as_ompl_state writes exactly RobotState.group_qpos, sub-space by sub-space."""
    from ramp_b200 import MotionPlanner, RobotState, load_robot_model, spaces

    rm = load_robot_model("zoo::" + cfg)
    planner = MotionPlanner.__new__(MotionPlanner)  # no GPU handle needed for the codec
    planner._robot_model, planner._group_name = rm, "arm"
    planner.space = spaces.make_space(rm, "arm")
    ospace = space_for(planner.space)
    rs = RobotState.from_named_state(rm, "arm", "home")
    vals = spaces.sample_uniform(planner.space, 1, np.random.default_rng(3))[0]
    rs.set_group_qpos("arm", vals)
    st = planner.as_ompl_state(rs, ospace, make_state(ospace))
    assert np.allclose(oc.from_ompl_state(ospace, st), rs.group_qpos("arm"), atol=1e-12)


@pytest.mark.gpu
def test_state_validity_callable_and_motion_validator(built):
    """The (state) -> bool callable equals is_state_valid on the decoded values, and the
    MotionValidator built over a stand-in ``ob`` module answers like the GPU validator."""
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    from ramp_b200 import MotionPlanner, RobotState, spaces, workloads
    from ramp_b200.motion_planner import make_ompl_motion_validator

    rm, rs = workloads.build("rrr_float")
    planner = MotionPlanner(rm, "arm")
    ref = RobotState.from_named_state(rm, "arm", "home")
    planner.set_reference_state(ref)
    ospace = space_for(planner.space)
    checker = planner.state_validity_checker(ref, ospace)
    rng = np.random.default_rng(5)
    samples = spaces.sample_uniform(planner.space, 64, rng, position_bound=1.0)
    want = planner.are_states_valid(samples)
    got = []
    for v in samples:
        st = make_state(ospace)
        oc.fill_ompl_state(ospace, st, v)
        got.append(checker(st))
    assert got == list(want) and 0 < sum(got) < len(got)

    class _MV:
        def __init__(self, si):
            self.si = si

    class _Ob:
        MotionValidator = _MV

    class _SI:
        def getStateSpace(self):  # noqa: N802
            return ospace

    mv = make_ompl_motion_validator(planner, _SI(), _Ob)
    for a, b in zip(samples[:16], samples[16:32]):
        sa, sb = make_state(ospace), make_state(ospace)
        oc.fill_ompl_state(ospace, sa, a)
        oc.fill_ompl_state(ospace, sb, b)
        assert mv.checkMotion(sa, sb) == planner.validator.checkMotion(a, b)
