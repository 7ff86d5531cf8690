"""The oracle's shape-shape signed distance against analytic geometry and independent solvers.

The reference pins no numeric collision result, so before the oracle judges
the GPU kernels it is itself checked here: closed-form cases derived by hand, an independent
convex-hull distance solver (SLSQP over barycentric weights), the exact box-box penetration
depth from the separating-axis theorem, and invariances (symmetry, rigid motion).
"""

import itertools

import numpy as np
import pytest
from scipy.optimize import minimize

from oracle import oracle
from ramp_b200.se3 import SE3, rpy_to_matrix
from ramp_b200.shapes import Box, Capsule, Convex, Cylinder, Ellipsoid, Sphere

I = SE3.Identity()


def T(x, y, z, rpy=(0.0, 0.0, 0.0)):
    return SE3(rpy_to_matrix(*rpy), [x, y, z])


def dist(a, ma, b, mb):
    return oracle.shape_distance(a, ma, b, mb)[0]


ANALYTIC = [
    (Sphere(0.5), I, Sphere(0.25), T(1, 0, 0), 0.25),
    (Sphere(0.5), I, Sphere(0.25), T(0.5, 0, 0), -0.25),
    (Box(1, 1, 1), I, Box(1, 1, 1), T(2, 0, 0), 1.0),
    (Box(1, 1, 1), I, Box(1, 1, 1), T(0.8, 0, 0), -0.2),
    (Box(1, 1, 1), I, Box(1, 1, 1), T(2, 0, 0, (0, 0, np.pi / 4)), 2 - 0.5 - np.sqrt(0.5)),
    (Box(1, 1, 1), I, Box(1, 1, 1), T(2, 2, 0), np.sqrt(2.0)),
    (Cylinder(0.5, 1), I, Cylinder(0.5, 1), T(2, 0, 0), 1.0),
    (Cylinder(0.5, 1), I, Cylinder(0.5, 1), T(0.9, 0, 0), -0.1),
    (Cylinder(0.5, 1), I, Cylinder(0.5, 1), T(0, 0, 0.7), -0.3),
    (Cylinder(0.5, 1), I, Cylinder(0.5, 1), T(0, 0, 1.7), 0.7),
    (Capsule(0.1, 0.4), I, Box(1, 1, 1), T(1, 0, 0), 0.4),
    (Capsule(0.1, 0.4), I, Box(1, 1, 1), T(0.55, 0, 0), -0.05),
    (Capsule(0.1, 0.4), I, Box(1, 1, 1), T(0.3, 0, 0), -0.3),
    (Capsule(0.1, 0.4), I, Capsule(0.1, 0.4), T(0.5, 0, 0), 0.3),
    (Capsule(0.1, 0.4), I, Capsule(0.1, 0.4), T(0, 0, 1.0), 0.4),
    (Capsule(0.1, 0.4), I, Capsule(0.2, 0.4), T(0.25, 0, 0, (np.pi / 2, 0, 0)), -0.05),
    (Sphere(0.1), I, Capsule(0.1, 0.4), T(0.5, 0, 0.1), 0.3),
    (Sphere(0.1), I, Box(1, 1, 1), T(1, 0, 0), 0.4),
    (Sphere(0.1), I, Box(1, 1, 1), T(0.1, 0, 0), -0.5),
    (Sphere(0.1), I, Box(1, 1, 1), T(1, 1, 1), np.sqrt(0.75) - 0.1),
    (Sphere(0.1), I, Cylinder(0.5, 1), T(1, 0, 0), 0.4),
    (Sphere(0.1), I, Cylinder(0.5, 1), T(0, 0, 1), 0.4),
    (Sphere(0.1), I, Cylinder(0.5, 1), T(0.2, 0, 0.1), -0.4),
    (Ellipsoid(0.5, 0.3, 0.2), I, Ellipsoid(0.5, 0.3, 0.2), T(2, 0, 0), 1.0),
    (Ellipsoid(0.5, 0.3, 0.2), I, Ellipsoid(0.5, 0.3, 0.2), T(0, 0, 0.3), -0.1),
    (Ellipsoid(0.5, 0.3, 0.2), I, Box(1, 1, 1), T(0, 1.0, 0), 0.2),
    (Cylinder(0.5, 1), I, Box(1, 1, 1), T(1.2, 0, 0), 0.2),
    (Cylinder(0.5, 1), I, Capsule(0.1, 0.4), T(0.8, 0, 0), 0.2),
]


@pytest.mark.parametrize("a,ma,b,mb,expected", ANALYTIC)
def test_analytic_distances(a, ma, b, mb, expected):
    # curved-vs-curved penetration converges through EPA's polytope: looser there
    tol = 2e-3 if expected < 0 and not isinstance(a, (Box, Sphere)) else 1e-6
    assert dist(a, ma, b, mb) == pytest.approx(expected, abs=tol)
    assert dist(b, mb, a, ma) == pytest.approx(expected, abs=tol)  # symmetry


def _hull_distance(pa, pb):
    """min |sum(l_i a_i) - sum(m_j b_j)| over the two probability simplices."""
    na, nb = len(pa), len(pb)

    def f(x):
        d = x[:na] @ pa - x[na:] @ pb
        return d @ d

    def g(x):
        d = x[:na] @ pa - x[na:] @ pb
        return np.concatenate([2 * pa @ d, -2 * pb @ d])

    cons = [{"type": "eq", "fun": lambda x: x[:na].sum() - 1}, {"type": "eq", "fun": lambda x: x[na:].sum() - 1}]
    x0 = np.concatenate([np.full(na, 1 / na), np.full(nb, 1 / nb)])
    r = minimize(f, x0, jac=g, bounds=[(0, 1)] * (na + nb), constraints=cons, method="SLSQP",
                 options={"ftol": 1e-14, "maxiter": 500})
    return np.sqrt(max(r.fun, 0.0))


def _random_pose(rng, spread):
    return T(*rng.uniform(-spread, spread, 3), rpy=tuple(rng.uniform(-np.pi, np.pi, 3)))


def test_convex_separated_distance_vs_independent_solver():
    rng = np.random.default_rng(0)
    checked = 0
    for _ in range(60):
        pa, pb = rng.normal(size=(12, 3)) * 0.3, rng.normal(size=(10, 3)) * 0.3
        ma, mb = _random_pose(rng, 0.2), _random_pose(rng, 1.2)
        wa = pa @ ma.rotation.T + ma.translation
        wb = pb @ mb.rotation.T + mb.translation
        ref = _hull_distance(wa, wb)
        got = dist(Convex(pa), ma, Convex(pb), mb)
        if ref > 1e-3:
            assert got == pytest.approx(ref, abs=2e-6 + 2e-6 * ref)
            checked += 1
        else:
            assert got <= 1e-6
    assert checked > 20


def _box_sat_signed(ha, ma, hb, mb):
    """Exact signed separation of two boxes: max over the 15 SAT axes of the projected gap."""
    ra, rb = ma.rotation, mb.rotation
    axes = [ra[:, i] for i in range(3)] + [rb[:, i] for i in range(3)]
    for i, j in itertools.product(range(3), range(3)):
        c = np.cross(ra[:, i], rb[:, j])
        if np.linalg.norm(c) > 1e-9:
            axes.append(c / np.linalg.norm(c))
    t = mb.translation - ma.translation
    best = -np.inf
    for ax in axes:
        gap = abs(t @ ax) - np.abs(ra.T @ ax) @ ha - np.abs(rb.T @ ax) @ hb
        best = max(best, gap)
    return best


def test_box_box_penetration_depth_is_sat_depth():
    """For two boxes the penetration depth is the least SAT overlap (faces of the Minkowski
    difference are face normals and edge-edge cross products)."""
    rng = np.random.default_rng(1)
    n_pen = 0
    for _ in range(200):
        a, b = Box(*rng.uniform(0.2, 1.0, 3)), Box(*rng.uniform(0.2, 1.0, 3))
        ma, mb = _random_pose(rng, 0.1), _random_pose(rng, 0.5)
        sat = _box_sat_signed(a.halfSide, ma, b.halfSide, mb)
        got = dist(a, ma, b, mb)
        if sat < -1e-6:
            assert got == pytest.approx(sat, abs=1e-6)
            n_pen += 1
        else:
            assert got >= -1e-9 and got >= sat - 1e-9  # SAT gap is a lower bound when separated
    assert n_pen > 50


def test_box_equals_its_vertex_hull():
    rng = np.random.default_rng(2)
    for _ in range(100):
        sides = rng.uniform(0.1, 1.0, 3)
        pts = np.array(list(itertools.product(*[(-s / 2, s / 2) for s in sides])))
        other = Capsule(0.1, 0.5)
        ma, mb = _random_pose(rng, 0.1), _random_pose(rng, 0.7)
        assert dist(Box(*sides), ma, other, mb) == pytest.approx(dist(Convex(pts), ma, other, mb), abs=1e-6)


def test_rigid_motion_invariance():
    rng = np.random.default_rng(3)
    shapes = [Sphere(0.2), Capsule(0.1, 0.4), Box(0.3, 0.2, 0.5), Cylinder(0.15, 0.4), Ellipsoid(0.3, 0.2, 0.1),
              Convex(rng.normal(size=(16, 3)) * 0.2)]
    for a, b in itertools.combinations_with_replacement(shapes, 2):
        ma, mb, g = _random_pose(rng, 0.3), _random_pose(rng, 0.6), _random_pose(rng, 2.0)
        d0 = dist(a, ma, b, mb)
        d1 = dist(a, g * ma, b, g * mb)
        tol = 5e-3 if d0 < 0 else 2e-6
        assert d1 == pytest.approx(d0, abs=tol)


def test_capsule_is_swept_segment():
    """coal reduces capsules/spheres to segment/point cores and sweeps the radius afterwards:
    distances between cores shift by the radii exactly."""
    rng = np.random.default_rng(4)
    for _ in range(50):
        ma, mb = _random_pose(rng, 0.1), _random_pose(rng, 1.0)
        box = Box(0.3, 0.4, 0.2)
        thin, fat = Capsule(0.01, 0.5), Capsule(0.11, 0.5)
        d_thin, d_fat = dist(thin, ma, box, mb), dist(fat, ma, box, mb)
        if d_thin > 0.0:
            assert d_fat == pytest.approx(d_thin - 0.1, abs=1e-6)


def test_deep_cylinder_box_penetration_with_coplanar_supports():
    """Two poses found by the 5M-configuration GPU sweep (profiles/r02_parity_sweep.jsonl): the flat
    cap of the cylinder against a flat box face gives EPA many support points in ONE plane.  The
    expected depths are the minimum over 2e6 directions (then Nelder-Mead) of the support function of
    the Minkowski difference; before the fix to epa_run's visibility test the oracle returned +0.107
    / +0.203 (and a different depth with the shapes swapped)."""
    cases = [([0.46854342343891586, -0.768924520186733, -0.4349967156724901, -0.7824039276186465, -0.4712425369778934, 0.1989564691339747,
               -0.8592710834130415, 0.935761938783602, 0.7472600161475853, 0.607594770963992, -0.26913019647622605, 0.9855978028821824],
              [-0.877776274907609, -0.3117564933656693, -0.36375362553763296, -0.7852042357784582, -0.337422411468121, -0.13667946244025087,
               0.9313779258944908, 0.8786426800345262, -0.3400807661666772, 0.9402800718321422, 0.014780361627331803, 0.976366464164009],
              -0.10728902115057082),
             ([0.9099034187656029, 0.17977561070692172, 0.373840202545925, 0.6463237340080641, 0.18709062628437767, 0.6264925373330422,
               -0.7566400682892823, 0.22672571581618772, -0.37023352738662696, 0.7584113825485769, 0.5364133699527809, -0.1124738068193623],
              [-0.9202227285901365, -0.2548877022044138, -0.29702253749006124, 0.6317938122956098, -0.386509836875086, 0.47226838433511265,
               0.7921948711488737, 0.24208032294738835, -0.06164637654202596, 0.8437978584167546, -0.5331087054579247, -0.12545676683813073],
              -0.1435206731861588)]
    cyl, box = Cylinder(0.1, 0.05), Box(0.1, 0.1, 0.3)
    for m1, m2, depth in cases:
        a, b = np.asarray(m1).reshape(3, 4), np.asarray(m2).reshape(3, 4)
        A, B = SE3(a[:, :3], a[:, 3]), SE3(b[:, :3], b[:, 3])
        d12, _ = oracle.shape_distance(cyl, A, box, B)
        d21, _ = oracle.shape_distance(box, B, cyl, A)
        assert abs(d12 - depth) < 2e-7 and abs(d21 - depth) < 2e-7, (d12, d21, depth)


def _support_value(geo, d):
    """max over the shape of d . x, for rows of local directions d."""
    name = type(geo).__name__
    if name == "Box":
        return (np.abs(d) * np.asarray(geo.halfSide)).sum(1)
    if name == "Cylinder":
        return geo.radius * np.linalg.norm(d[:, :2], axis=1) + geo.halfLength * np.abs(d[:, 2])
    if name == "Capsule":
        return geo.radius * np.linalg.norm(d, axis=1) + geo.halfLength * np.abs(d[:, 2])
    if name == "Convex":
        return (d @ np.asarray(geo.points).T).max(1)
    raise AssertionError(name)


def _signed_distance_by_direction_search(a, ma, b, mb, rng):
    """Signed distance of two convex shapes as -min over unit n of the support function of A - B:
    positive = separation, negative = penetration depth.  Independent of GJK / EPA."""
    Ra, ta, Rb, tb = ma.rotation, ma.translation, mb.rotation, mb.translation

    def h(n):
        return _support_value(a, n @ Ra) + n @ ta + _support_value(b, (-n) @ Rb) - n @ tb

    n = rng.normal(size=(200000, 3))
    n /= np.linalg.norm(n, axis=1, keepdims=True)
    hv = h(n)
    f = lambda x: h((x / np.linalg.norm(x))[None])[0]  # noqa: E731
    best = min((minimize(f, n[j], method="Nelder-Mead", options=dict(xatol=1e-12, fatol=1e-14, maxiter=4000)) for j in np.argsort(hv)[:4]),
               key=lambda r: r.fun)
    return -best.fun


def test_signed_distances_against_a_direction_search():
    """Boxes, cylinders, capsules and 40-vertex hulls in random and in axis-aligned relative poses (flat faces
    parallel: the case that produces coplanar support points), separated and overlapping, in both
    argument orders, against the minimum of the support function of the Minkowski difference."""
    rng = np.random.default_rng(42)
    def hull():  # a 40-vertex polytope, the kind of shape the stand-in links are
        p = rng.normal(size=(40, 3))
        return Convex(p / np.linalg.norm(p, axis=1, keepdims=True) * rng.uniform(0.05, 0.2, size=(1, 3)))

    makers = [lambda: Box(*rng.uniform(0.05, 0.3, 3)), lambda: Cylinder(rng.uniform(0.03, 0.12), rng.uniform(0.05, 0.3)),
              lambda: Capsule(rng.uniform(0.03, 0.1), rng.uniform(0.05, 0.3)), hull]

    def rot(k):
        if k % 2:  # axis-aligned up to a signed permutation
            r = np.eye(3)[:, rng.permutation(3)] * rng.choice([-1.0, 1.0], 3)
            if np.linalg.det(r) < 0:
                r[:, 0] *= -1
            return r
        return rpy_to_matrix(*rng.uniform(-np.pi, np.pi, 3))

    worst, n_pen = 0.0, 0
    for k in range(48):
        a, b = makers[k % 4](), makers[(k // 4) % 4]()
        ma = SE3(rot(k), np.zeros(3))
        mb = SE3(rot(k // 2), rng.normal(size=3) * rng.uniform(0.02, 0.3))
        truth = _signed_distance_by_direction_search(a, ma, b, mb, rng)
        d12, _ = oracle.shape_distance(a, ma, b, mb)
        d21, _ = oracle.shape_distance(b, mb, a, ma)
        tol = 2e-6 + 2e-5 * abs(truth)
        assert abs(d12 - truth) < tol and abs(d21 - truth) < tol, (k, type(a).__name__, type(b).__name__, truth, d12, d21)
        worst = max(worst, abs(d12 - truth), abs(d21 - truth))
        n_pen += truth < 0
    print("worst deviation", worst, "overlapping cases", n_pen)
    assert 10 <= n_pen <= 38
