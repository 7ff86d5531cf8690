"""ctypes front-end of the CPU oracle (oracle/ramp_oracle.c).

TEST INFRASTRUCTURE ONLY -- see the header of ramp_oracle.c.  Importable from tests/,
__graft_entry__.smoke() and bench.py's CPU-baseline legs; never from ramp_b200/.
"""

from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

from ramp_b200.desc import IkParams, ModelDesc, QMap, RampGeom, RampModelDesc, RampQMap, RampSpaceDesc, SpaceDesc

HERE = Path(__file__).resolve().parent
LIB_PATH = HERE / "libramp_oracle.so"

MODE_STOP_AT_FIRST = 1
MODE_DEPTH = 2
MODE_EXACT = 4


def build(force: bool = False) -> Path:
    src = HERE / "ramp_oracle.c"
    if force or not LIB_PATH.exists() or LIB_PATH.stat().st_mtime < src.stat().st_mtime:
        subprocess.run(["make", "-C", str(HERE), "-s"], check=True)
    return LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(str(LIB_PATH))
        _lib.orc_shape_distance.restype = C.c_double
    return _lib


def _p(a, t):
    return None if a is None else a.ctypes.data_as(C.POINTER(t))


def set_params(gjk_tol=1e-6, gjk_max_iter=128, epa_tol=1e-6, epa_max_iter=64):
    lib().orc_set_params(C.c_double(gjk_tol), C.c_int(gjk_max_iter), C.c_double(epa_tol), C.c_int(epa_max_iter))


def fk(desc: ModelDesc, q: np.ndarray) -> np.ndarray:
    q = np.ascontiguousarray(np.atleast_2d(q), dtype=np.float64)
    n, nj = len(q), desc.struct.n_joints
    out = np.zeros((n, nj, 3, 4))
    rc = lib().orc_fk_batch(C.byref(desc.struct), _p(q, C.c_double), C.c_int64(n), C.c_int(q.shape[1]), _p(out, C.c_double))
    assert rc == 0
    return out


def placements(desc: ModelDesc, q: np.ndarray) -> np.ndarray:
    q = np.ascontiguousarray(np.atleast_2d(q), dtype=np.float64)
    n = len(q)
    out = np.zeros((n, desc.struct.n_geoms, 3, 4))
    rc = lib().orc_placements_batch(C.byref(desc.struct), _p(q, C.c_double), C.c_int64(n), C.c_int(q.shape[1]), _p(out, C.c_double))
    assert rc == 0
    return out


def check(desc: ModelDesc, q: np.ndarray, *, mode: int = MODE_DEPTH | MODE_EXACT, active=None,
          want_pair_dist: bool = True, nthreads: int = 0) -> dict:
    """computeCollisions restated.  Returns hit, first_pair, min_dist, pair_dist, iters."""
    q = np.ascontiguousarray(np.atleast_2d(q), dtype=np.float64)
    n, npairs = len(q), desc.struct.n_pairs
    hit = np.zeros(n, dtype=np.uint8)
    first = np.zeros(n, dtype=np.int32)
    mind = np.zeros(n)
    pd = np.zeros((n, npairs)) if want_pair_dist else None
    iters = np.zeros(n, dtype=np.int64)
    act = None if active is None else np.ascontiguousarray(active, dtype=np.uint8)
    rc = lib().orc_check_batch(C.byref(desc.struct), _p(act, C.c_uint8), _p(q, C.c_double), C.c_int64(n),
                               C.c_int(q.shape[1]), C.c_int(mode), _p(hit, C.c_uint8), _p(first, C.c_int32),
                               _p(mind, C.c_double), _p(pd, C.c_double), _p(iters, C.c_int64), C.c_int(nthreads))
    assert rc == 0
    return {"hit": hit.astype(bool), "first_pair": first, "min_dist": mind, "pair_dist": pd, "iters": iters}


def check_f32(desc: ModelDesc, q: np.ndarray, nthreads: int = 0) -> np.ndarray:
    """Boolean, stop-at-first check on a float32 batch: the timed CPU baseline."""
    assert q.dtype == np.float32 and q.flags.c_contiguous
    hit = np.zeros(len(q), dtype=np.uint8)
    rc = lib().orc_check_batch_f32(C.byref(desc.struct), None, _p(q, C.c_float), C.c_int64(len(q)),
                                   C.c_int(q.shape[1]), C.c_int(MODE_STOP_AT_FIRST), _p(hit, C.c_uint8),
                                   C.c_int(nthreads))
    assert rc == 0
    return hit


def flops(desc: ModelDesc, q: np.ndarray, stop_at_first: bool = False) -> np.ndarray:
    """Algorithmic FP32 flops per configuration (SURVEY 8d formula; roofline numerator).
    ``stop_at_first``: a configuration's pair list ends at its first colliding pair, as in the
    reference's boolean call (and in the timed GPU call)."""
    q = np.ascontiguousarray(np.atleast_2d(q), dtype=np.float64)
    out = np.zeros(len(q))
    rc = lib().orc_flops_batch(C.byref(desc.struct), _p(q, C.c_double), C.c_int64(len(q)), C.c_int(q.shape[1]),
                               C.c_int(int(stop_at_first)), _p(out, C.c_double))
    assert rc == 0
    return out


def ik(desc: ModelDesc, qmap: QMap, params: IkParams, seeds: np.ndarray, targets: np.ndarray) -> dict:
    """FP64 statement of ramp_ik_batch: ``seeds`` [n, n_act] actuated values, ``targets`` [n, 3, 4] or
    [3, 4] world poses of the tool frame.  Returns q, err, ok, iters."""
    seeds = np.ascontiguousarray(np.atleast_2d(seeds), dtype=np.float64)
    targets = np.ascontiguousarray(targets, dtype=np.float64)
    one = targets.ndim == 2
    n = len(seeds)
    q = np.zeros((n, qmap.n_act))
    err = np.zeros(n)
    ok = np.zeros(n, dtype=np.uint8)
    iters = np.zeros(n, dtype=np.int32)
    rc = lib().orc_ik_batch(C.byref(desc.struct), C.byref(qmap.struct), C.byref(params.struct), _p(seeds, C.c_double), C.c_int64(n),
                            C.c_int(seeds.shape[1]), _p(targets.reshape(-1), C.c_double), C.c_int(int(one)), _p(q, C.c_double),
                            _p(err, C.c_double), _p(ok, C.c_uint8), _p(iters, C.c_int32))
    assert rc == 0
    return {"q": q, "err": err, "ok": ok.astype(bool), "iters": iters}


def num_threads() -> int:
    return int(lib().orc_num_threads())


def shape_distance(geom1, m1, geom2, m2) -> tuple[float, int]:
    """Signed distance of two placed GeometryObject-like shapes (placement = world pose)."""
    from ramp_b200.desc import fill_geoms

    class _G:
        def __init__(self, geometry, placement):
            self.geometry, self.placement, self.parentJoint = geometry, placement, 0

    arr, verts = fill_geoms([_G(geom1, m1), _G(geom2, m2)])
    a = np.ascontiguousarray(m1.rt34().reshape(-1))
    b = np.ascontiguousarray(m2.rt34().reshape(-1))
    it = C.c_int(0)
    v = np.ascontiguousarray(verts.reshape(-1)) if len(verts) else np.zeros(3)
    d = lib().orc_shape_distance(C.byref(arr[0]), _p(a, C.c_double), C.byref(arr[1]), _p(b, C.c_double),
                                 _p(v, C.c_double), C.byref(it))
    return float(d), it.value


def qmap_apply(qmap: QMap, act: np.ndarray) -> np.ndarray:
    act = np.ascontiguousarray(np.atleast_2d(act), dtype=np.float64)
    out = np.zeros((len(act), qmap.nq))
    lib().orc_qmap_apply(C.byref(qmap.struct), C.c_int(qmap.nq), _p(act, C.c_double), C.c_int64(len(act)),
                         C.c_int(act.shape[1]), _p(out, C.c_double))
    return out


def edge_nd(space: SpaceDesc, q0, q1) -> int:
    a = np.ascontiguousarray(q0, dtype=np.float64)
    b = np.ascontiguousarray(q1, dtype=np.float64)
    return int(lib().orc_edge_nd(C.byref(space.struct), _p(a, C.c_double), _p(b, C.c_double)))


def edge_interpolate(space: SpaceDesc, q0, q1, t: float) -> np.ndarray:
    a = np.ascontiguousarray(q0, dtype=np.float64)
    b = np.ascontiguousarray(q1, dtype=np.float64)
    out = np.zeros_like(a)
    lib().orc_edge_interpolate(C.byref(space.struct), _p(a, C.c_double), _p(b, C.c_double), C.c_double(t), _p(out, C.c_double))
    return out


def satisfies_bounds(space: SpaceDesc, q) -> bool:
    a = np.ascontiguousarray(q, dtype=np.float64)
    return bool(lib().orc_satisfies_bounds(C.byref(space.struct), _p(a, C.c_double)))


def check_edges(desc: ModelDesc, qmap: QMap, space: SpaceDesc, q0, q1, active=None, nthreads: int = 0) -> dict:
    q0 = np.ascontiguousarray(np.atleast_2d(q0), dtype=np.float64)
    q1 = np.ascontiguousarray(np.atleast_2d(q1), dtype=np.float64)
    n = len(q0)
    valid = np.zeros(n, dtype=np.uint8)
    first = np.zeros(n, dtype=np.int32)
    nd = np.zeros(n, dtype=np.int32)
    act = None if active is None else np.ascontiguousarray(active, dtype=np.uint8)
    rc = lib().orc_check_edges(C.byref(desc.struct), _p(act, C.c_uint8), C.byref(qmap.struct), C.byref(space.struct),
                               _p(q0, C.c_double), _p(q1, C.c_double), C.c_int64(n), C.c_int(q0.shape[1]),
                               _p(valid, C.c_uint8), _p(first, C.c_int32), _p(nd, C.c_int32), C.c_int(nthreads))
    assert rc == 0
    return {"valid": valid.astype(bool), "first_invalid": first, "nd": nd}


__all__ = [n for n in dir() if not n.startswith("_")] + ["RampGeom", "RampModelDesc", "RampQMap", "RampSpaceDesc"]
