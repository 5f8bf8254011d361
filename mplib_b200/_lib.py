"""ctypes binding of libmpb.so (include/mpb.h).  Fails loudly: there is no CPU compute path."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from .scene import MpbSceneDesc, MpbStats, SceneDesc


class MpbSpaceDesc(C.Structure):
    """``mpb_space_desc``: the compound planning space of ``OMPLPlannerTpl::build_state_space``."""

    _fields_ = [("dim", C.c_int32), ("lo", C.POINTER(C.c_float)), ("hi", C.POINTER(C.c_float)),
                ("seg", C.POINTER(C.c_float)), ("so2", C.POINTER(C.c_uint8))]


class MpbRrtParams(C.Structure):
    _fields_ = [("range", C.c_float), ("max_rounds", C.c_int32), ("time_limit_s", C.c_float), ("tree_capacity", C.c_int32),
                ("max_path", C.c_int32), ("seed", C.c_uint64), ("check_every", C.c_int32), ("connect_steps", C.c_int32),
                ("n_slots", C.c_int32), ("max_rounds_per_query", C.c_int32)]


class MpbRrtStats(C.Structure):
    _fields_ = [("rounds", C.c_int64), ("grow_calls", C.c_int64), ("states_checked", C.c_int64), ("n_solved", C.c_int64),
                ("kernel_launches", C.c_int64), ("gpu_ms", C.c_float), ("host_ms_total", C.c_float)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


RRT_ACTIVE, RRT_SOLVED, RRT_CAPACITY, RRT_INVALID_GOAL, RRT_PATH_TOO_LONG = range(5)


class Space:
    """Host arrays of a planning space + the ctypes struct pointing into them."""

    def __init__(self, lo, hi, so2=None, resolution=0.01):
        self.lo = np.ascontiguousarray(np.asarray(lo, dtype=np.float32))
        self.hi = np.ascontiguousarray(np.asarray(hi, dtype=np.float32))
        self.dim = int(self.lo.shape[0])
        self.so2 = np.ascontiguousarray(np.zeros(self.dim, np.uint8) if so2 is None else np.asarray(so2, dtype=np.uint8))
        # longest valid segment per subspace: resolution * maximum extent (pi for SO(2))  [upstream OMPL]
        ext = np.where(self.so2 != 0, np.pi, self.hi.astype(np.float64) - self.lo.astype(np.float64))
        self.seg64 = resolution * ext
        self.seg = np.ascontiguousarray(self.seg64.astype(np.float32))
        self.struct = MpbSpaceDesc(self.dim, self.lo.ctypes.data_as(C.POINTER(C.c_float)), self.hi.ctypes.data_as(C.POINTER(C.c_float)),
                                   self.seg.ctypes.data_as(C.POINTER(C.c_float)), self.so2.ctypes.data_as(C.POINTER(C.c_uint8)))

MPB_HOST, MPB_DEVICE = 0, 1
_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("MPB_LIB_PATH") or os.path.join(_HERE, "libmpb.so")  # override: debugging builds only

# every symbol include/mpb.h declares (tests check that the library exports all of them)
SYMBOLS = [
    "mpb_last_error", "mpb_device_count", "mpb_scene_create", "mpb_scene_destroy", "mpb_scene_set_q_base",
    "mpb_scene_set_pairs", "mpb_fk_batch", "mpb_collide_batch", "mpb_scene_info", "mpb_fp32_peak",
    "mpb_check_motion_batch", "mpb_rrtconnect_batch", "mpb_distance_batch", "mpb_ik_clik_batch",
    "mpb_link_jacobian_batch", "mpb_collide_batch_hits", "mpb_scene_poll", "mpb_scene_set_async_timing",
    "mpb_scene_update_objects", "mpb_fk_batch_noalloc", "mpb_debug_cell_table",
]

_lib = None


class MpbError(RuntimeError):
    pass


# The pybind11 host module (csrc/pymp_native.cpp) carries the per-call path when it is built: numpy buffers in, C-ABI
# call with the GIL released, result words unpacked in C++.  The ctypes binding below stays as the fallback BINDING
# (never a compute path) and serves the pointer-level calls.
try:
    if os.environ.get("MPB_NO_NATIVE"):
        raise ImportError("disabled by MPB_NO_NATIVE")
    from . import _native
except ImportError:
    _native = None


def _nat(f, *a, **k):
    """Call into the native module with the binding's exception types (MpbError for non-argument failures)."""
    try:
        return f(*a, **k)
    except ValueError:
        raise
    except RuntimeError as e:
        if isinstance(e, MpbError) or str(e).startswith("State dimension"):
            raise
        raise MpbError(str(e)) from None


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `make` (or __graft_entry__.build()). "
            "mplib_b200 has no CPU fallback for the collision path."
        )
    lib = C.CDLL(LIB_PATH)
    lib.mpb_last_error.restype = C.c_char_p
    lib.mpb_device_count.restype = C.c_int
    lib.mpb_scene_create.argtypes = [C.POINTER(MpbSceneDesc), C.c_int, C.POINTER(C.c_void_p)]
    lib.mpb_scene_destroy.argtypes = [C.c_void_p]
    lib.mpb_scene_destroy.restype = None
    lib.mpb_scene_set_q_base.argtypes = [C.c_void_p, C.c_void_p, C.c_int32]
    lib.mpb_scene_set_pairs.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p]
    lib.mpb_fk_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int, C.c_void_p]
    lib.mpb_debug_cell_table.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_int64, C.POINTER(C.c_int64)]
    lib.mpb_fk_batch_noalloc.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]
    lib.mpb_collide_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p,
                                      C.POINTER(MpbStats), C.c_int, C.c_void_p]
    lib.mpb_collide_batch_hits.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p,
                                           C.POINTER(MpbStats), C.c_int, C.c_void_p]
    lib.mpb_scene_poll.argtypes = [C.c_void_p, C.c_int, C.POINTER(MpbStats), C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
    lib.mpb_scene_set_async_timing.argtypes = [C.c_void_p, C.c_int]
    lib.mpb_scene_update_objects.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.mpb_scene_info.argtypes = [C.c_void_p] + [C.POINTER(C.c_int32)] * 5
    lib.mpb_fp32_peak.argtypes = [C.c_int, C.POINTER(C.c_float)]
    lib.mpb_distance_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.POINTER(MpbStats), C.c_int,
                                       C.c_void_p]
    lib.mpb_check_motion_batch.argtypes = [C.c_void_p, C.POINTER(MpbSpaceDesc), C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p,
                                           C.POINTER(MpbStats), C.c_int, C.c_void_p]
    lib.mpb_rrtconnect_batch.argtypes = [C.c_void_p, C.POINTER(MpbSpaceDesc), C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32,
                                         C.c_int64, C.POINTER(MpbRrtParams), C.c_void_p, C.c_void_p, C.c_void_p,
                                         C.POINTER(MpbRrtStats)]
    lib.mpb_ik_clik_batch.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_double, C.c_int32,
                                      C.c_double, C.c_double, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.mpb_link_jacobian_batch.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_void_p]
    _lib = lib
    return lib


def _check(rc):
    if rc != 0:
        msg = load().mpb_last_error().decode("utf-8", "replace")
        if rc == -1:
            raise ValueError(msg)
        raise MpbError(f"libmpb error {rc}: {msg}")


def _ptr(x):
    """numpy array -> host pointer; torch tensor -> data_ptr; int -> as is."""
    if x is None:
        return None
    if isinstance(x, np.ndarray):
        return x.ctypes.data
    if hasattr(x, "data_ptr"):
        return x.data_ptr()
    return int(x)


class DeviceScene:
    """Owner of one compiled ``mpb_scene`` (device image + queues) on one GPU."""

    def __init__(self, desc: SceneDesc, device: int = 0):
        lib = load()
        self.desc = desc
        self.device = device
        self._n = None
        if _native is not None:
            self._n = _nat(_native.Scene, {k: getattr(desc, k) for k in desc._ARRAYS}, device)
            self._h = C.c_void_p(self._n.handle)  # the same handle serves the pointer-level ctypes calls
        else:
            h = C.c_void_p()
            _check(lib.mpb_scene_create(C.byref(desc.as_struct()), device, C.byref(h)))
            self._h = h
        self.n_pairs = desc.n_pairs
        self.pair_words = (desc.n_pairs + 31) // 32
        self.state_dim = desc.state_dim
        self.n_links = desc.n_links

    def close(self):
        if getattr(self, "_n", None) is not None:
            self._n.close()
            self._n = None
            self._h = None
        elif getattr(self, "_h", None) and _lib is not None:  # _lib is None during interpreter shutdown
            _lib.mpb_scene_destroy(self._h)
            self._h = None

    __del__ = close

    def set_q_base(self, q_base):
        q = np.ascontiguousarray(np.asarray(q_base, dtype=np.float64))
        if self._n is not None:
            _nat(self._n.set_q_base, q)
        else:
            _check(load().mpb_scene_set_q_base(self._h, q.ctypes.data, q.shape[0]))
        self.desc = self.desc.with_q_base(q)

    def set_pairs(self, pair_a, pair_b, pair_kinds=()):
        a = np.ascontiguousarray(np.asarray(pair_a, dtype=np.int32))
        b = np.ascontiguousarray(np.asarray(pair_b, dtype=np.int32))
        if self._n is not None:
            _nat(self._n.set_pairs, a, b)
        else:
            _check(load().mpb_scene_set_pairs(self._h, a.shape[0], a.ctypes.data, b.ctypes.data))
        self.desc = self.desc.with_pairs(a, b, pair_kinds)
        self.n_pairs = int(a.shape[0])
        self.pair_words = (self.n_pairs + 31) // 32

    def update_objects(self, idx, obj_local=None, obj_param=None):
        """In-place pose (row-major 3x4 per object) / primitive-parameter update of compiled objects."""
        idx = np.ascontiguousarray(np.asarray(idx, dtype=np.int32).reshape(-1))
        loc = None if obj_local is None else np.ascontiguousarray(np.asarray(obj_local, dtype=np.float64).reshape(-1, 12))
        par = None if obj_param is None else np.ascontiguousarray(np.asarray(obj_param, dtype=np.float64).reshape(-1, 4))
        if self._n is not None:
            _nat(self._n.update_objects, idx, loc, par)
        else:
            _check(load().mpb_scene_update_objects(self._h, idx.shape[0], _ptr(idx), _ptr(loc), _ptr(par)))
        self.desc = self.desc.with_objects(idx, loc, par)

    def poll(self, wait=True):
        """Harvest finished device-pointer calls -> (summed stats dict, calls completed, calls failed).  Raises MpbError
        when a harvested call overflowed the work queues."""
        if self._n is not None:
            st, done, failed = _nat(self._n.poll, bool(wait))
            return st, int(done), int(failed)
        st = MpbStats()
        done, failed = C.c_int64(0), C.c_int64(0)
        _check(load().mpb_scene_poll(self._h, 1 if wait else 0, C.byref(st), C.byref(done), C.byref(failed)))
        return st.as_dict(), int(done.value), int(failed.value)

    def set_async_timing(self, on=True):
        _check(load().mpb_scene_set_async_timing(self._h, 1 if on else 0))

    def collide_hits(self, q, hit_cap=None, stats=False):
        """q [n, D] host -> (verdict bool[n], hits int64[k, 2] = (configuration, pair) sorted[, stats]): the sparse form
        of ``collide(want_pairs=True)``."""
        q = np.ascontiguousarray(np.asarray(q, dtype=np.float32))
        if q.ndim != 2 or q.shape[1] != self.state_dim:
            raise RuntimeError(f"State dimension is not correct: expected [n,{self.state_dim}], got {q.shape}")
        n = q.shape[0]
        if self._n is not None:
            verdict, h, st = _nat(self._n.collide_hits, q, int(hit_cap) if hit_cap is not None else 0)
            return (verdict, h, st) if stats else (verdict, h)
        cap = int(hit_cap) if hit_cap is not None else max(1024, 4 * n)
        st = MpbStats()
        while True:
            vb = np.zeros((n + 31) // 32, dtype=np.uint32)
            hits = np.zeros((max(cap, 1), 2), dtype=np.uint32)
            nh = np.zeros(1, dtype=np.uint32)
            _check(load().mpb_collide_batch_hits(self._h, _ptr(q), n, _ptr(vb), _ptr(hits), cap, _ptr(nh), C.byref(st), MPB_HOST, None))
            if int(nh[0]) <= cap or hit_cap is not None:
                break
            cap = int(nh[0])  # the list was truncated: repeat with the reported size
        verdict = np.unpackbits(vb.view(np.uint8), bitorder="little")[:n].astype(bool)
        h = hits[: min(int(nh[0]), cap)].astype(np.int64)
        h = h[np.lexsort((h[:, 1], h[:, 0]))]
        return (verdict, h, st.as_dict()) if stats else (verdict, h)

    # -- host-buffer entry points -------------------------------------------------------------
    def collide(self, q, want_pairs=True, stats=False):
        """q: [n, state_dim] (any float dtype, host).  -> (verdict bool[n], pair_hit bool[n,P] | None[, stats])"""
        q = np.ascontiguousarray(np.asarray(q, dtype=np.float32))
        if q.ndim != 2 or q.shape[1] != self.state_dim:
            raise RuntimeError(f"State dimension is not correct: expected [n,{self.state_dim}], got {q.shape}")
        n = q.shape[0]
        if self._n is not None:
            verdict, pairs, st = _nat(self._n.collide, q, bool(want_pairs), bool(stats))
            return (verdict, pairs, st) if stats else (verdict, pairs)
        vb = np.zeros((n + 31) // 32, dtype=np.uint32)
        pb = np.zeros((n, max(self.pair_words, 1)), dtype=np.uint32) if want_pairs else None
        st = MpbStats()
        _check(load().mpb_collide_batch(self._h, _ptr(q), n, _ptr(vb), _ptr(pb), C.byref(st), MPB_HOST, None))
        verdict = np.unpackbits(vb.view(np.uint8), bitorder="little")[:n].astype(bool)
        pairs = None
        if want_pairs:
            pairs = np.unpackbits(pb.view(np.uint8), axis=1, bitorder="little")[:, : self.n_pairs].astype(bool)
        return (verdict, pairs, st.as_dict()) if stats else (verdict, pairs)

    def collide_raw(self, q_ptr, n, verdict_ptr, pair_ptr=None, space=MPB_HOST, stream=None, stats=None):
        """Pointer-level call (pinned host or device memory); returns the stats dict."""
        st = MpbStats()
        _check(load().mpb_collide_batch(self._h, _ptr(q_ptr), int(n), _ptr(verdict_ptr), _ptr(pair_ptr), C.byref(st),
                                        space, _ptr(stream)))
        return st.as_dict()

    def collide_hits_raw(self, q_ptr, n, verdict_ptr, hits_ptr, hit_cap, n_hits_ptr, space=MPB_HOST, stream=None):
        """Pointer-level mpb_collide_batch_hits (pinned host or device memory); returns the stats dict."""
        st = MpbStats()
        _check(load().mpb_collide_batch_hits(self._h, _ptr(q_ptr), int(n), _ptr(verdict_ptr), _ptr(hits_ptr), int(hit_cap),
                                             _ptr(n_hits_ptr), C.byref(st), space, _ptr(stream)))
        return st.as_dict()

    def fk(self, q):
        q = np.ascontiguousarray(np.asarray(q, dtype=np.float32))
        if q.ndim != 2 or q.shape[1] != self.state_dim:
            raise RuntimeError(f"State dimension is not correct: expected [n,{self.state_dim}], got {q.shape}")
        if self._n is not None:
            return _nat(self._n.fk, q)
        out = np.zeros((q.shape[0], self.n_links, 7), dtype=np.float32)
        _check(load().mpb_fk_batch(self._h, _ptr(q), q.shape[0], _ptr(out), MPB_HOST, None))
        return out

    def distance(self, q, stats=False):
        """q [n, D] -> (min_distance float64[n] (-1 where an admitted pair collides, float64 max if no pair),
        argmin pair int32[n])"""
        q = np.ascontiguousarray(np.asarray(q, dtype=np.float32))
        if q.ndim != 2 or q.shape[1] != self.state_dim:
            raise RuntimeError(f"State dimension is not correct: expected [n,{self.state_dim}], got {q.shape}")
        n = q.shape[0]
        if self._n is not None:
            out, am, st = _nat(self._n.distance, q)
            return (out, am, st) if stats else (out, am)
        md = np.zeros(n, dtype=np.float32)
        am = np.zeros(n, dtype=np.int32)
        st = MpbStats()
        _check(load().mpb_distance_batch(self._h, _ptr(q), n, _ptr(md), _ptr(am), C.byref(st), MPB_HOST, None))
        out = md.astype(np.float64)
        out[am < 0] = np.finfo(np.float64).max  # WorldDistanceResult::min_distance starts at max (src/planning_world.h:30-41)
        return (out, am, st.as_dict()) if stats else (out, am)

    def check_motion(self, space: Space, a, b, stats=False):
        """Batched DiscreteMotionValidator::checkMotion: a, b [E, D] -> valid bool[E]."""
        a = np.ascontiguousarray(np.asarray(a, dtype=np.float32))
        b = np.ascontiguousarray(np.asarray(b, dtype=np.float32))
        if a.ndim != 2 or a.shape != b.shape or a.shape[1] != self.state_dim:
            raise RuntimeError(f"State dimension is not correct: expected two [E,{self.state_dim}] arrays")
        E = a.shape[0]
        vb = np.zeros((E + 31) // 32, dtype=np.uint32)
        st = MpbStats()
        _check(load().mpb_check_motion_batch(self._h, C.byref(space.struct), _ptr(a), _ptr(b), E, _ptr(vb), C.byref(st), MPB_HOST, None))
        valid = np.unpackbits(vb.view(np.uint8), bitorder="little")[:E].astype(bool)
        return (valid, st.as_dict()) if stats else valid

    def rrtconnect(self, space: Space, starts, goals, n_goals=None, range_=0.1, max_rounds=4096, time_limit=0.0,
                   tree_capacity=2048, max_path=1024, seed=0, check_every=16, connect_steps=8, n_slots=0, max_rounds_per_query=0):
        """Lock-step RRTConnect for Q queries.  starts [Q, D]; goals [Q, G, D] (or [Q, D]); n_goals [Q].
        -> (status int32[Q], list of float64 paths ([n_i, D], empty if unsolved), stats dict)"""
        starts = np.ascontiguousarray(np.asarray(starts, dtype=np.float32))
        goals = np.asarray(goals, dtype=np.float32)
        if goals.ndim == 2:
            goals = goals[:, None, :]
        goals = np.ascontiguousarray(goals)
        Q, D = starts.shape
        if D != self.state_dim or goals.shape[0] != Q or goals.shape[2] != D:
            raise RuntimeError("Length of start state and goal state should be equal")
        G = goals.shape[1]
        ng = np.ascontiguousarray(np.full(Q, G, np.int32) if n_goals is None else np.asarray(n_goals, dtype=np.int32))
        prm = MpbRrtParams(float(range_), int(max_rounds), float(time_limit), int(tree_capacity), int(max_path), int(seed), int(check_every), int(connect_steps), int(n_slots), int(max_rounds_per_query))
        status = np.zeros(Q, dtype=np.int32)
        plen = np.zeros(Q, dtype=np.int32)
        paths = np.empty((Q, max_path, D), dtype=np.float32)  # rows past path_len stay unwritten
        st = MpbRrtStats()
        _check(load().mpb_rrtconnect_batch(self._h, C.byref(space.struct), _ptr(starts), _ptr(goals), _ptr(ng), G, Q, C.byref(prm),
                                           _ptr(status), _ptr(plen), _ptr(paths), C.byref(st)))
        out = [paths[i, : plen[i]].astype(np.float64) for i in range(Q)]
        return status, out, st.as_dict()

    def ik_clik(self, link, poses, q_init, mask=None, eps=1e-5, max_iter=1000, dt=1e-1, damp=1e-12):
        """Batched computeIKCLIK: poses [n, 7], q_init [n, D] -> (q [n, D] float64, success bool[n], err [n, 6])."""
        poses = np.ascontiguousarray(np.asarray(poses, dtype=np.float64).reshape(-1, 7))
        q0 = np.ascontiguousarray(np.asarray(q_init, dtype=np.float64))
        if q0.ndim != 2 or q0.shape[1] != self.state_dim or q0.shape[0] != poses.shape[0]:
            raise RuntimeError("Qpos user2pinocchio failed")
        n = q0.shape[0]
        m = None
        if mask is not None and len(mask):
            m = np.ascontiguousarray(np.asarray(mask, dtype=np.uint8))
            if m.shape[0] != self.state_dim:
                raise RuntimeError("mask length does not match the number of joints")
        q = np.zeros((n, self.state_dim), dtype=np.float64)
        ok = np.zeros(n, dtype=np.uint8)
        err = np.zeros((n, 6), dtype=np.float64)
        _check(load().mpb_ik_clik_batch(self._h, int(link), _ptr(poses), _ptr(q0), _ptr(m), n, float(eps), int(max_iter), float(dt),
                                        float(damp), _ptr(q), _ptr(ok), _ptr(err)))
        return q, ok.astype(bool), err

    def link_jacobian(self, link, q, local=False):
        """Batched computeFullJacobian + getLinkJacobian: q [n, D] -> (J [n, 6, D], link placement [n, 3, 4]), float64."""
        q = np.ascontiguousarray(np.asarray(q, dtype=np.float64))
        if q.ndim != 2 or q.shape[1] != self.state_dim:
            raise RuntimeError("Qpos user2pinocchio failed")
        n = q.shape[0]
        J = np.zeros((n, 6, self.state_dim), dtype=np.float64)
        T = np.zeros((n, 3, 4), dtype=np.float64)
        _check(load().mpb_link_jacobian_batch(self._h, int(link), _ptr(q), n, 1 if local else 0, _ptr(J), _ptr(T)))
        return J, T

    def fk_raw(self, q_ptr, n, poses_ptr, space=MPB_HOST, stream=None):
        _check(load().mpb_fk_batch(self._h, _ptr(q_ptr), int(n), _ptr(poses_ptr), space, _ptr(stream)))

    def fk_noalloc_raw(self, q_dev_ptr, n, poses_dev_ptr, stream=None):
        """mpb_fk_batch_noalloc: device pointers, no allocation inside the call, returns after the enqueue."""
        _check(load().mpb_fk_batch_noalloc(self._h, _ptr(q_dev_ptr), int(n), _ptr(poses_dev_ptr), _ptr(stream)))


def debug_cell_table(verts):
    """Support cell table of a hull (diagnostics): -> (first uint16[6*G*G+1], candidates uint8[...])."""
    v = np.ascontiguousarray(np.asarray(verts, dtype=np.float64))
    nb = C.c_int64(0)
    _check(load().mpb_debug_cell_table(v.ctypes.data, v.shape[0], None, 0, C.byref(nb)))
    buf = np.zeros(nb.value, dtype=np.uint8)
    _check(load().mpb_debug_cell_table(v.ctypes.data, v.shape[0], buf.ctypes.data, buf.size, C.byref(nb)))
    n_first = 6 * 8 * 8 + 1
    return buf[:2 * n_first].view(np.uint16).copy(), buf[2 * n_first:].copy()


def fp32_peak_tflops(device=0):
    v = C.c_float(0)
    _check(load().mpb_fp32_peak(device, C.byref(v)))
    return float(v.value)
