"""ctypes binding of libramp_b200.so and the ``GpuModel`` handle.

This is the only module that talks to the CUDA library.  It fails loudly when the library
is not built or no CUDA device is usable -- there is no CPU fallback anywhere in ramp_b200.
PyTorch supplies device memory and streams (``tensor.data_ptr()``, current stream); all
arithmetic happens inside the library's kernels.
"""

from __future__ import annotations

import ctypes as C
import os
import weakref
from pathlib import Path

import numpy as np
import torch

from ramp_b200 import desc as D
from ramp_b200.exceptions import RampGpuError, RampLibraryMissing
from ramp_b200.kinematics import GeometryModel, Model

LIB_PATH = Path(__file__).resolve().parent / "libramp_b200.so"

_lib = None


def load_library():
    """dlopen libramp_b200.so (built in-tree by __graft_entry__.build) and type its entry points."""
    global _lib
    if _lib is not None:
        return _lib
    path = Path(os.environ.get("RAMP_B200_LIB", LIB_PATH))
    if not path.exists():
        raise RampLibraryMissing(
            f"{path} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'`. "
            "ramp_b200 has no CPU fallback.")
    lib = C.CDLL(str(path))
    vp, i32, i64, u32 = C.c_void_p, C.c_int32, C.c_int64, C.c_uint32
    sig = {
        "ramp_abi_version": (C.c_int, []),
        "ramp_last_error": (C.c_char_p, []),
        "ramp_launch_count": (i64, []),
        "ramp_model_create": (C.c_int, [C.POINTER(D.RampModelDesc), C.c_int, C.POINTER(vp)]),
        "ramp_model_set_scene": (C.c_int, [vp, C.POINTER(D.RampGeom), i32, C.POINTER(i32), i32, C.POINTER(C.c_double), i32]),
        "ramp_model_set_qmap": (C.c_int, [vp, C.POINTER(D.RampQMap)]),
        "ramp_model_set_pair_active": (C.c_int, [vp, i32, i32]),
        "ramp_model_destroy": (None, [vp]),
        "ramp_model_trim": (C.c_int, [vp]),
        "ramp_model_info": (C.c_int, [vp, C.POINTER(i32)]),
        "ramp_model_counters": (C.c_int, [vp, C.POINTER(C.c_uint64), i32]),
        "ramp_fk_batch": (C.c_int, [vp, vp, i64, i32, u32, vp, vp]),
        "ramp_placements_batch": (C.c_int, [vp, vp, i64, i32, u32, vp, vp]),
        "ramp_check_batch": (C.c_int, [vp, vp, i64, i32, u32, vp, vp, vp, vp, vp]),
        "ramp_check_batch_noalloc": (C.c_int, [vp, vp, i64, i32, u32, vp, vp, vp, vp]),
        "ramp_pack_flags": (C.c_int, [vp, vp, i64, vp, vp]),
        "ramp_ik_batch": (C.c_int, [vp, vp, i64, i32, vp, i32, C.POINTER(D.RampIkParams), vp, vp, vp, vp, vp]),
        "ramp_distance_batch": (C.c_int, [vp, vp, i64, i32, u32, C.POINTER(i32), i32, vp, vp]),
        "ramp_check_edges": (C.c_int, [vp, vp, vp, i64, i32, C.POINTER(D.RampSpaceDesc), vp, vp, vp, vp]),
        "ramp_check_edges_host": (C.c_int, [vp, vp, vp, i64, i32, C.POINTER(D.RampSpaceDesc), vp, vp, vp]),
        "ramp_check_host": (C.c_int, [vp, vp, i64, i32, u32, vp, vp, vp, vp]),
        "ramp_host_alloc": (vp, [C.c_size_t]),
        "ramp_host_free": (None, [vp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)  # AttributeError here = header and library disagree
        fn.restype, fn.argtypes = res, args
    if lib.ramp_abi_version() != D.ABI_VERSION:
        raise RampLibraryMissing(f"{path}: ABI {lib.ramp_abi_version()} != {D.ABI_VERSION}; rebuild")
    _lib = lib
    return lib


def _check(rc: int, what: str):
    if rc != 0:
        msg = load_library().ramp_last_error().decode(errors="replace")
        raise RampGpuError(f"{what} failed ({rc}): {msg}")


def launch_count() -> int:
    return int(load_library().ramp_launch_count())


def pinned_empty(shape, dtype=np.float32) -> np.ndarray:
    """A numpy array in page-locked host memory (ramp_host_alloc): the host-buffer calls copy
    straight from / to such arrays instead of staging them.  Freed when the array is collected."""
    lib = load_library()
    dtype = np.dtype(dtype)
    nbytes = int(np.prod(shape, dtype=np.int64)) * dtype.itemsize
    ptr = lib.ramp_host_alloc(nbytes)
    if not ptr:
        raise RampGpuError("ramp_host_alloc failed: " + lib.ramp_last_error().decode(errors="replace"))
    buf = (C.c_byte * max(nbytes, 1)).from_address(ptr)
    weakref.finalize(buf, lib.ramp_host_free, ptr)
    return np.frombuffer(buf, dtype=dtype, count=nbytes // dtype.itemsize).reshape(shape)


def _stream_ptr(device: torch.device) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def _ptr(t: torch.Tensor | None) -> int | None:
    return None if t is None else t.data_ptr()


class GpuModel:
    """Device-resident robot + scene tables and the batched entry points over them."""

    def __init__(self, model: Model, geom_model: GeometryModel, device: int | None = None):
        self.lib = load_library()
        if not torch.cuda.is_available():
            raise RampGpuError("no CUDA device: ramp_b200 has no CPU fallback")
        self.device_index = torch.cuda.current_device() if device is None else int(device)
        self.device = torch.device("cuda", self.device_index)
        self.model = model
        self.n_robot_geoms = geom_model.ngeoms
        self.n_robot_pairs = len(geom_model.collisionPairs)
        self._robot_pairs = [(p.first, p.second) for p in geom_model.collisionPairs]
        self.desc = D.ModelDesc(model, geom_model)
        self.handle = C.c_void_p()
        _check(self.lib.ramp_model_create(C.byref(self.desc.struct), self.device_index, C.byref(self.handle)),
               "ramp_model_create")
        self._scene_sig = (self.n_robot_geoms, tuple(self._robot_pairs))
        self.qmap: D.QMap | None = None
        self._keep = []

    def __del__(self):
        h = getattr(self, "handle", None)
        if h:
            self.lib.ramp_model_destroy(h)
            self.handle = None

    # ------------------------------------------------------------------ model state
    def info(self) -> dict:
        out = (C.c_int32 * 6)()
        _check(self.lib.ramp_model_info(self.handle, out), "ramp_model_info")
        return dict(zip(("nq", "n_joints", "n_geoms", "n_pairs", "n_act", "device"), list(out)))

    def counters(self, reset: bool = False) -> dict:
        """GJK work items the queue kernel consumed and its launches since creation / the last reset."""
        out = (C.c_uint64 * 2)()
        _check(self.lib.ramp_model_counters(self.handle, out, int(reset)), "ramp_model_counters")
        return {"gjk_items": int(out[0]), "gjk_queue_launches": int(out[1])}

    def sync(self, geom_model: GeometryModel) -> None:
        """Mirror mutations of the host collision model (RobotState.add_object appends
        universe objects and pairs; disable_collision_pair removes pairs) onto the device."""
        pairs = [(p.first, p.second) for p in geom_model.collisionPairs]
        sig = (geom_model.ngeoms, tuple(pairs))
        if sig == self._scene_sig:
            return
        tail_objs = geom_model.geometryObjects[self.n_robot_geoms:]
        head_ok = pairs[: self.n_robot_pairs] == self._robot_pairs and all(o.parentJoint == 0 for o in tail_objs)
        if head_ok:
            arr, verts = D.fill_geoms(tail_objs)
            extra = np.ascontiguousarray(pairs[self.n_robot_pairs:], dtype=np.int32).reshape(-1, 2)
            _check(self.lib.ramp_model_set_scene(
                self.handle, C.cast(arr, C.POINTER(D.RampGeom)), len(tail_objs),
                extra.ctypes.data_as(C.POINTER(C.c_int32)), len(extra),
                verts.ctypes.data_as(C.POINTER(C.c_double)), len(verts)), "ramp_model_set_scene")
        else:  # robot pairs changed: rebuild the handle from scratch
            self.lib.ramp_model_destroy(self.handle)
            self.n_robot_geoms, self.n_robot_pairs, self._robot_pairs = geom_model.ngeoms, len(pairs), pairs
            self.desc = D.ModelDesc(self.model, geom_model)
            self.handle = C.c_void_p()
            _check(self.lib.ramp_model_create(C.byref(self.desc.struct), self.device_index, C.byref(self.handle)),
                   "ramp_model_create")
            if self.qmap is not None:
                _check(self.lib.ramp_model_set_qmap(self.handle, C.byref(self.qmap.struct)), "ramp_model_set_qmap")
        self._scene_sig = sig

    def trim(self) -> None:
        """Hand the scratch pool's cached memory back to the device (ramp_model_trim)."""
        _check(self.lib.ramp_model_trim(self.handle), "ramp_model_trim")

    def set_qmap(self, qmap: D.QMap) -> None:
        self.qmap = qmap
        _check(self.lib.ramp_model_set_qmap(self.handle, C.byref(qmap.struct)), "ramp_model_set_qmap")

    def set_pair_active(self, pair_index: int, active: bool) -> None:
        _check(self.lib.ramp_model_set_pair_active(self.handle, int(pair_index), int(active)), "ramp_model_set_pair_active")

    # ------------------------------------------------------------------ batched calls
    def _q(self, q) -> torch.Tensor:
        if not isinstance(q, torch.Tensor):
            q = torch.as_tensor(np.asarray(q, dtype=np.float32))
        q = q.to(device=self.device, dtype=torch.float32)
        if q.dim() == 1:
            q = q[None]
        return q.contiguous()

    def forward_kinematics(self, q, *, actuated: bool = False) -> torch.Tensor:
        """q[N, nq] -> oMi[N, n_joints, 3, 4] (ramp_fk_batch)."""
        q = self._q(q)
        nj = self.model.njoints - 1
        out = torch.empty((q.shape[0], nj, 3, 4), dtype=torch.float32, device=self.device)
        _check(self.lib.ramp_fk_batch(self.handle, q.data_ptr(), q.shape[0], q.shape[1],
                                      D.Q_ACTUATED if actuated else 0, out.data_ptr(), _stream_ptr(self.device)),
               "ramp_fk_batch")
        return out

    def placements(self, q, *, actuated: bool = False) -> torch.Tensor:
        """q[N, nq] -> oMg[N, n_geoms, 3, 4] (ramp_placements_batch)."""
        q = self._q(q)
        ng = self.info()["n_geoms"]
        out = torch.empty((q.shape[0], ng, 3, 4), dtype=torch.float32, device=self.device)
        _check(self.lib.ramp_placements_batch(self.handle, q.data_ptr(), q.shape[0], q.shape[1],
                                              D.Q_ACTUATED if actuated else 0, out.data_ptr(), _stream_ptr(self.device)),
               "ramp_placements_batch")
        return out

    def check(self, q, *, actuated: bool = False, first_pair: bool = False, pair_mask: bool = False,
              all_pairs: bool = False, min_dist: bool = False, out_hit: torch.Tensor | None = None) -> dict:
        """Batched computeCollisions (ramp_check_batch).  Returns ``hit`` (bool[N]) and, on
        request, ``first_pair`` (int32[N], -1 = none), ``pair_mask`` (bool[N, P]) and ``min_dist``
        (float32[N], smallest signed distance over the active pairs; a much slower kernel)."""
        q = self._q(q)
        n = q.shape[0]
        npairs = self.info()["n_pairs"]
        hit = out_hit if out_hit is not None else torch.empty(n, dtype=torch.uint8, device=self.device)
        first = torch.empty(n, dtype=torch.int32, device=self.device) if first_pair else None
        words = (npairs + 31) // 32
        mask = torch.zeros((n, max(words, 1)), dtype=torch.int32, device=self.device) if pair_mask else None
        flags = (D.Q_ACTUATED if actuated else 0) | (D.ALL_PAIRS if all_pairs or pair_mask else 0)
        dmin = torch.empty(n, dtype=torch.float32, device=self.device) if min_dist else None
        _check(self.lib.ramp_check_batch(self.handle, q.data_ptr(), n, q.shape[1], flags, hit.data_ptr(), _ptr(first),
                                         _ptr(dmin), _ptr(mask), _stream_ptr(self.device)), "ramp_check_batch")
        res = {"hit": hit.bool()}
        if dmin is not None:
            res["min_dist"] = dmin
        if first is not None:
            res["first_pair"] = first
        if mask is not None:
            bits = (mask[:, :, None] >> torch.arange(32, device=self.device, dtype=torch.int32)) & 1
            res["pair_mask"] = bits.reshape(n, -1)[:, :npairs].bool()
        return res

    def distance(self, q, pair_idx=None, *, actuated: bool = False) -> torch.Tensor:
        """Signed distance of the selected pairs (all pairs when ``pair_idx`` is None):
        float32[N, n_sel], positive separation / negative penetration depth (ramp_distance_batch)."""
        q = self._q(q)
        n = q.shape[0]
        if pair_idx is None:
            sel, n_sel = None, self.info()["n_pairs"]
        else:
            arr = np.ascontiguousarray(pair_idx, dtype=np.int32)
            sel, n_sel = arr.ctypes.data_as(C.POINTER(C.c_int32)), len(arr)
        out = torch.empty((n, n_sel), dtype=torch.float32, device=self.device)
        _check(self.lib.ramp_distance_batch(self.handle, q.data_ptr(), n, q.shape[1], D.Q_ACTUATED if actuated else 0,
                                            sel, n_sel, out.data_ptr(), _stream_ptr(self.device)), "ramp_distance_batch")
        return out

    def clearance(self, q, *, actuated: bool = False) -> torch.Tensor:
        """Smallest signed distance over the enabled pairs, float32[N] (+inf without pairs): the
        quantity the reference's PathClearanceObjective integrates (src/ramp/motion_planner.py:165-206)."""
        q = self._q(q)
        if self.info()["n_pairs"] == 0:
            return torch.full((q.shape[0],), float("inf"), dtype=torch.float32, device=self.device)
        # the d_min_dist output of ramp_check_batch: pairs whose capsule bounds cannot lower the
        # minimum skip GJK / EPA (a tenth of the work of the full distance matrix on an arm)
        return self.check(q, actuated=actuated, min_dist=True)["min_dist"]

    def check_raw(self, q: torch.Tensor, hit: torch.Tensor, flags: int = 0) -> None:
        """Zero-overhead variant for benchmarks: caller-owned contiguous tensors, no allocations."""
        _check(self.lib.ramp_check_batch(self.handle, q.data_ptr(), q.shape[0], q.shape[1], flags, hit.data_ptr(),
                                         None, None, None, _stream_ptr(self.device)), "ramp_check_batch")

    def pack_flags(self, hit: torch.Tensor, words: torch.Tensor) -> None:
        """uint8 flags [n] -> int32 bit words [ceil(n / 32)] on the device (ramp_pack_flags)."""
        _check(self.lib.ramp_pack_flags(self.handle, hit.data_ptr(), hit.shape[0], words.data_ptr(), _stream_ptr(self.device)),
               "ramp_pack_flags")

    def ik_batch(self, seeds, targets, params: D.IkParams) -> dict:
        """Batched differential IK (ramp_ik_batch) on the actuated values of the group the qmap
        describes: ``seeds`` [n, n_act]; ``targets`` [n, 3, 4] or one [3, 4] world pose of the tool
        frame.  Returns ``q`` [n, n_act], ``err`` [n], ``ok`` bool[n], ``iters`` int32[n]."""
        seeds = self._q(seeds)
        n = seeds.shape[0]
        tg = torch.as_tensor(np.asarray(targets, dtype=np.float32) if not isinstance(targets, torch.Tensor) else targets)
        tg = tg.to(device=self.device, dtype=torch.float32).contiguous()
        one = tg.dim() == 2
        assert tg.shape[-2:] == (3, 4) and (one or tg.shape[0] == n)
        n_act = self.info()["n_act"]
        q = torch.empty((n, n_act), dtype=torch.float32, device=self.device)
        err = torch.empty(n, dtype=torch.float32, device=self.device)
        ok = torch.empty(n, dtype=torch.uint8, device=self.device)
        iters = torch.empty(n, dtype=torch.int32, device=self.device)
        _check(self.lib.ramp_ik_batch(self.handle, seeds.data_ptr(), n, seeds.shape[1], tg.data_ptr(), int(one),
                                      C.byref(params.struct), q.data_ptr(), err.data_ptr(), ok.data_ptr(), iters.data_ptr(),
                                      _stream_ptr(self.device)), "ramp_ik_batch")
        return {"q": q, "err": err, "ok": ok.bool(), "iters": iters}

    def check_edges(self, q0, q1, space: D.SpaceDesc) -> dict:
        """Batched DiscreteMotionValidator (ramp_check_edges) over actuated edge end points."""
        q0, q1 = self._q(q0), self._q(q1)
        assert q0.shape == q1.shape
        n = q0.shape[0]
        valid = torch.empty(n, dtype=torch.uint8, device=self.device)
        first = torch.empty(n, dtype=torch.int32, device=self.device)
        nd = torch.empty(n, dtype=torch.int32, device=self.device)
        _check(self.lib.ramp_check_edges(self.handle, q0.data_ptr(), q1.data_ptr(), n, q0.shape[1],
                                         C.byref(space.struct), valid.data_ptr(), first.data_ptr(), nd.data_ptr(),
                                         _stream_ptr(self.device)), "ramp_check_edges")
        return {"valid": valid.bool(), "first_invalid": first, "nd": nd}

    def check_edges_host(self, q0: np.ndarray, q1: np.ndarray, space: D.SpaceDesc) -> dict:
        """ramp_check_edges_host: numpy end points in, numpy verdicts out, one library call."""
        q0 = np.ascontiguousarray(np.atleast_2d(q0), dtype=np.float32)
        q1 = np.ascontiguousarray(np.atleast_2d(q1), dtype=np.float32)
        n = len(q0)
        valid = np.zeros(n, dtype=np.uint8)
        first = np.zeros(n, dtype=np.int32)
        nd = np.zeros(n, dtype=np.int32)
        _check(self.lib.ramp_check_edges_host(self.handle, q0.ctypes.data, q1.ctypes.data, n, q0.shape[1],
                                              C.byref(space.struct), valid.ctypes.data, first.ctypes.data, nd.ctypes.data),
               "ramp_check_edges_host")
        return {"valid": valid.astype(bool), "first_invalid": first, "nd": nd}

    def check_host_into(self, q: np.ndarray, hit: np.ndarray, *, actuated: bool = False) -> None:
        """ramp_check_host on caller-owned buffers, nothing allocated or converted: q float32
        C-contiguous [n, ld], hit uint8 [n].  Page-locked arrays (``pinned_empty``) are copied
        without staging."""
        if q.dtype != np.float32 or not q.flags.c_contiguous or q.ndim != 2:
            raise ValueError("q must be a C-contiguous float32 [n, ld] array")
        if hit.dtype != np.uint8 or not hit.flags.c_contiguous or hit.size < len(q):
            raise ValueError("hit must be a C-contiguous uint8 array of at least n elements")
        _check(self.lib.ramp_check_host(self.handle, q.ctypes.data, len(q), q.shape[1], D.Q_ACTUATED if actuated else 0,
                                        hit.ctypes.data, None, None, None), "ramp_check_host")

    def check_host(self, q: np.ndarray, *, actuated: bool = False, first_pair: bool = False,
                   pair_mask: bool = False, min_dist: bool = False) -> dict:
        """Host-buffer call (ramp_check_host): numpy in, numpy out, copies inside the library."""
        q = np.ascontiguousarray(np.atleast_2d(q), dtype=np.float32)
        n = len(q)
        npairs = self.info()["n_pairs"]
        words = max((npairs + 31) // 32, 1)
        hit = np.zeros(n, dtype=np.uint8)
        first = np.zeros(n, dtype=np.int32) if first_pair else None
        mask = np.zeros((n, words), dtype=np.uint32) if pair_mask else None
        flags = (D.Q_ACTUATED if actuated else 0) | (D.ALL_PAIRS if pair_mask else 0)
        dmin = np.zeros(n, dtype=np.float32) if min_dist else None
        vp = lambda a: None if a is None else a.ctypes.data  # noqa: E731
        _check(self.lib.ramp_check_host(self.handle, q.ctypes.data, n, q.shape[1], flags, hit.ctypes.data, vp(first),
                                        vp(dmin), vp(mask)), "ramp_check_host")
        res = {"hit": hit.astype(bool)}
        if dmin is not None:
            res["min_dist"] = dmin
        if first is not None:
            res["first_pair"] = first
        if mask is not None:
            bits = (mask[:, :, None] >> np.arange(32, dtype=np.uint32)) & 1
            res["pair_mask"] = bits.reshape(n, -1)[:, :npairs].astype(bool)
        return res
