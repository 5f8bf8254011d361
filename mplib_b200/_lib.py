"""ctypes binding of libmpb.so (include/mpb.h).  Fails loudly: there is no CPU compute path."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from .scene import MpbSceneDesc, MpbStats, SceneDesc

MPB_HOST, MPB_DEVICE = 0, 1
_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libmpb.so")

# every symbol include/mpb.h declares (tests check that the library exports all of them)
SYMBOLS = [
    "mpb_last_error", "mpb_device_count", "mpb_scene_create", "mpb_scene_destroy", "mpb_scene_set_q_base",
    "mpb_scene_set_pairs", "mpb_fk_batch", "mpb_collide_batch", "mpb_scene_info", "mpb_fp32_peak",
]

_lib = None


class MpbError(RuntimeError):
    pass


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
    lib.mpb_collide_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p,
                                      C.POINTER(MpbStats), C.c_int, C.c_void_p]
    lib.mpb_scene_info.argtypes = [C.c_void_p] + [C.POINTER(C.c_int32)] * 5
    lib.mpb_fp32_peak.argtypes = [C.c_int, C.POINTER(C.c_float)]
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
        h = C.c_void_p()
        _check(lib.mpb_scene_create(C.byref(desc.as_struct()), device, C.byref(h)))
        self._h = h
        self.n_pairs = desc.n_pairs
        self.pair_words = (desc.n_pairs + 31) // 32
        self.state_dim = desc.state_dim
        self.n_links = desc.n_links

    def close(self):
        if getattr(self, "_h", None) and _lib is not None:  # _lib is None during interpreter shutdown
            _lib.mpb_scene_destroy(self._h)
            self._h = None

    __del__ = close

    def set_q_base(self, q_base):
        q = np.ascontiguousarray(np.asarray(q_base, dtype=np.float64))
        _check(load().mpb_scene_set_q_base(self._h, q.ctypes.data, q.shape[0]))
        self.desc = self.desc.with_q_base(q)

    def set_pairs(self, pair_a, pair_b, pair_kinds=()):
        a = np.ascontiguousarray(np.asarray(pair_a, dtype=np.int32))
        b = np.ascontiguousarray(np.asarray(pair_b, dtype=np.int32))
        _check(load().mpb_scene_set_pairs(self._h, a.shape[0], a.ctypes.data, b.ctypes.data))
        self.desc = self.desc.with_pairs(a, b, pair_kinds)
        self.n_pairs = int(a.shape[0])
        self.pair_words = (self.n_pairs + 31) // 32

    # -- host-buffer entry points -------------------------------------------------------------
    def collide(self, q, want_pairs=True, stats=False):
        """q: [n, state_dim] (any float dtype, host).  -> (verdict bool[n], pair_hit bool[n,P] | None[, stats])"""
        q = np.ascontiguousarray(np.asarray(q, dtype=np.float32))
        if q.ndim != 2 or q.shape[1] != self.state_dim:
            raise RuntimeError(f"State dimension is not correct: expected [n,{self.state_dim}], got {q.shape}")
        n = q.shape[0]
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

    def fk(self, q):
        q = np.ascontiguousarray(np.asarray(q, dtype=np.float32))
        if q.ndim != 2 or q.shape[1] != self.state_dim:
            raise RuntimeError(f"State dimension is not correct: expected [n,{self.state_dim}], got {q.shape}")
        out = np.zeros((q.shape[0], self.n_links, 7), dtype=np.float32)
        _check(load().mpb_fk_batch(self._h, _ptr(q), q.shape[0], _ptr(out), MPB_HOST, None))
        return out

    def fk_raw(self, q_ptr, n, poses_ptr, space=MPB_HOST, stream=None):
        _check(load().mpb_fk_batch(self._h, _ptr(q_ptr), int(n), _ptr(poses_ptr), space, _ptr(stream)))


def fp32_peak_tflops(device=0):
    v = C.c_float(0)
    _check(load().mpb_fp32_peak(device, C.byref(v)))
    return float(v.value)
