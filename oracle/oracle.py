"""ctypes wrapper of the CPU oracle (oracle/oracle.c).  TEST INFRASTRUCTURE ONLY -- see the header of
oracle.c.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may import this module."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "liboracle.so")
_lib = None


def build(force=False):
    src = os.path.join(_HERE, "oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        os.makedirs(os.path.dirname(_SO), exist_ok=True)
        subprocess.check_call(["gcc", "-O3", "-fopenmp", "-fPIC", "-shared", "-o", _SO, src, "-lm"])
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
        _lib.orc_object_distance.restype = C.c_double
    return _lib


def _states(desc, states):
    s = np.ascontiguousarray(np.asarray(states, dtype=np.float64))
    if s.ndim == 1:
        s = s.reshape(1, -1)
    assert s.shape[1] == desc.state_dim, (s.shape, desc.state_dim)
    return s


def fk_batch(desc, states):
    s = _states(desc, states)
    out = np.zeros((s.shape[0], desc.n_links, 7), dtype=np.float64)
    lib().orc_fk_batch(C.byref(desc.as_struct()), C.c_void_p(s.ctypes.data), C.c_int64(s.shape[0]), C.c_void_p(out.ctypes.data))
    return out


def object_poses(desc, states):
    s = _states(desc, states)
    out = np.zeros((s.shape[0], desc.n_objects, 12), dtype=np.float64)
    lib().orc_object_poses(C.byref(desc.as_struct()), C.c_void_p(s.ctypes.data), C.c_int64(s.shape[0]), C.c_void_p(out.ctypes.data))
    return out.reshape(s.shape[0], desc.n_objects, 3, 4)


class Accel:
    """Hierarchies of the TIMED baseline leg (orc_accel_create): AABB trees over mesh triangles / occupied voxels and
    hull adjacency for the edge-walking support.  Holds the description it was built from."""

    def __init__(self, desc):
        self.desc = desc
        lib().orc_accel_create.restype = C.c_void_p
        self.h = C.c_void_p(lib().orc_accel_create(C.byref(desc.as_struct())))

    def __del__(self):
        if getattr(self, "h", None) and _lib is not None:
            _lib.orc_accel_destroy.argtypes = [C.c_void_p]
            _lib.orc_accel_destroy(self.h)
            self.h = None


def collide_batch(desc, states, return_iters=False, threads=1, accel=None):
    """Exhaustive checker (accel=None) or, with ``accel=Accel(desc)``, the hierarchical leg that bench.py times."""
    s = _states(desc, states)
    n = s.shape[0]
    ph = np.zeros((n, max(desc.n_pairs, 1)), dtype=np.uint8)
    vd = np.zeros(n, dtype=np.uint8)
    it = C.c_int64(0)
    if accel is not None:
        assert accel.desc is desc, "the accelerator belongs to another scene description"
        lib().orc_collide_batch_accel(C.byref(desc.as_struct()), accel.h, C.c_void_p(s.ctypes.data), C.c_int64(n),
                                      C.c_void_p(ph.ctypes.data), C.c_void_p(vd.ctypes.data), C.byref(it), C.c_int(int(threads)))
    else:
        lib().orc_collide_batch_mt(C.byref(desc.as_struct()), C.c_void_p(s.ctypes.data), C.c_int64(n), C.c_void_p(ph.ctypes.data),
                                   C.c_void_p(vd.ctypes.data), C.byref(it), C.c_int(int(threads)))
    ph = ph[:, : desc.n_pairs].astype(bool)
    if return_iters:
        return vd.astype(bool), ph, it.value
    return vd.astype(bool), ph


def pair_signed_distance(desc, state, pair):
    s = _states(desc, state)
    out = C.c_double(0)
    lib().orc_pair_signed_distance(C.byref(desc.as_struct()), C.c_void_p(s.ctypes.data), C.c_int(pair), C.byref(out))
    return out.value


def sat_pair(desc, state, pair):
    s = _states(desc, state)
    out = C.c_double(0)
    rc = lib().orc_sat_pair(C.byref(desc.as_struct()), C.c_void_p(s.ctypes.data), C.c_int(pair), C.byref(out))
    if rc != 0:
        raise ValueError("orc_sat_pair: pair is not convex/convex")
    return out.value


def sat_batch(desc, states, item_state, item_pair, threads=8):
    s = _states(desc, states)
    ist = np.ascontiguousarray(np.asarray(item_state, dtype=np.int64))
    ip = np.ascontiguousarray(np.asarray(item_pair, dtype=np.int32))
    out = np.zeros(ist.shape[0], dtype=np.float64)
    rc = lib().orc_sat_batch_mt(C.byref(desc.as_struct()), C.c_void_p(s.ctypes.data), C.c_int64(ist.shape[0]),
                                C.c_void_p(ist.ctypes.data), C.c_void_p(ip.ctypes.data), C.c_void_p(out.ctypes.data),
                                C.c_int(int(threads)))
    if rc != 0:
        raise ValueError("orc_sat_batch_mt: a pair is not convex/convex")
    return out


def distance_batch(desc, states, threads=1, return_pairs=False):
    """min over the admitted pairs of ::fcl::distance (-1: some pair touches / intersects) + arg-min pair; mesh and
    octree pairs by exhaustive leaf minimum.  return_pairs: also the [n, n_pairs] matrix of pair distances."""
    s = _states(desc, states)
    n = s.shape[0]
    md = np.zeros(n, dtype=np.float64)
    am = np.zeros(n, dtype=np.int32)
    pd = np.zeros((n, max(desc.n_pairs, 1)), dtype=np.float64) if return_pairs else None
    lib().orc_distance_batch_ex(C.byref(desc.as_struct()), C.c_void_p(s.ctypes.data), C.c_int64(n), C.c_void_p(md.ctypes.data),
                                C.c_void_p(am.ctypes.data), C.c_void_p(pd.ctypes.data if return_pairs else None), C.c_int(int(threads)))
    if return_pairs:
        return md, am, pd[:, : desc.n_pairs]
    return md, am


def tri_tri_distance(A, B, closed=True):
    """Distance of two triangles ([3, 3] each): closed form (vertex-face + edge-edge) or the GJK the oracle's leaves use."""
    A = np.ascontiguousarray(np.asarray(A, dtype=np.float64).reshape(9))
    B = np.ascontiguousarray(np.asarray(B, dtype=np.float64).reshape(9))
    f = lib().orc_tri_tri_distance_closed if closed else lib().orc_tri_tri_distance_gjk
    f.restype = C.c_double
    return float(f(C.c_void_p(A.ctypes.data), C.c_void_p(B.ctypes.data)))


def tri_tri_intersect(A, B, independent=False):
    """SAT of FCL / PQP (default) or the independent edge-piercing + zero-distance-candidate test."""
    A = np.ascontiguousarray(np.asarray(A, dtype=np.float64).reshape(9))
    B = np.ascontiguousarray(np.asarray(B, dtype=np.float64).reshape(9))
    if independent:
        return bool(lib().orc_tri_tri_intersect_independent(C.c_void_p(A.ctypes.data), C.c_void_p(B.ctypes.data), C.c_double(1e-12)))
    return bool(lib().orc_tri_tri_intersect(C.c_void_p(A.ctypes.data), C.c_void_p(B.ctypes.data)))


def _T12(T):
    return np.ascontiguousarray(np.asarray(T, dtype=np.float64)[:3, :4].reshape(12))


def object_collide(desc, a, Ta, b, Tb):
    ta, tb = _T12(Ta), _T12(Tb)
    return bool(lib().orc_object_collide(C.byref(desc.as_struct()), C.c_int(a), C.c_void_p(ta.ctypes.data), C.c_int(b),
                                         C.c_void_p(tb.ctypes.data)))


def object_distance(desc, a, Ta, b, Tb):
    ta, tb = _T12(Ta), _T12(Tb)
    return float(lib().orc_object_distance(C.byref(desc.as_struct()), C.c_int(a), C.c_void_p(ta.ctypes.data), C.c_int(b),
                                           C.c_void_p(tb.ctypes.data)))


def _space_args(lo, hi, so2, resolution=0.01):
    lo = np.ascontiguousarray(np.asarray(lo, dtype=np.float64))
    hi = np.ascontiguousarray(np.asarray(hi, dtype=np.float64))
    so2 = np.ascontiguousarray(np.zeros(lo.shape[0], np.uint8) if so2 is None else np.asarray(so2, dtype=np.uint8))
    seg = np.ascontiguousarray(resolution * np.where(so2 != 0, np.pi, hi - lo))
    return lo, hi, seg, so2


def check_motion_batch(desc, lo, hi, a, b, so2=None, threads=8):
    """DiscreteMotionValidator::checkMotion per edge -> (valid bool[E], states checked)."""
    lo, hi, seg, so2 = _space_args(lo, hi, so2)
    a, b = _states(desc, a), _states(desc, b)
    E = a.shape[0]
    valid = np.zeros(E, dtype=np.uint8)
    ns = C.c_int64(0)
    rc = lib().orc_check_motion_batch(C.byref(desc.as_struct()), C.c_int(desc.state_dim), C.c_void_p(seg.ctypes.data),
                                      C.c_void_p(so2.ctypes.data), C.c_void_p(a.ctypes.data), C.c_void_p(b.ctypes.data), C.c_int64(E),
                                      C.c_void_p(valid.ctypes.data), C.byref(ns), C.c_int(int(threads)))
    assert rc == 0
    return valid.astype(bool), ns.value


def rrtconnect_batch(desc, lo, hi, starts, goals, n_goals=None, so2=None, range_=0.1, max_grow_calls=1 << 20, time_limit=1.0, seed=0,
                     max_path=4096, threads=1, accel=None):
    """CPU RRTConnect restatement for Q queries (one per thread at a time).
    -> (status int32[Q], list of paths, grow calls, state checks)"""
    lo, hi, seg, so2 = _space_args(lo, hi, so2)
    starts = _states(desc, starts)
    goals = np.asarray(goals, dtype=np.float64)
    if goals.ndim == 2:
        goals = goals[:, None, :]
    goals = np.ascontiguousarray(goals)
    Q, D = starts.shape
    G = goals.shape[1]
    ng = np.ascontiguousarray(np.full(Q, G, np.int32) if n_goals is None else np.asarray(n_goals, dtype=np.int32))
    status = np.zeros(Q, dtype=np.int32)
    plen = np.zeros(Q, dtype=np.int32)
    paths = np.zeros((Q, max_path, D), dtype=np.float64)
    gc, sc = C.c_int64(0), C.c_int64(0)
    lib().orc_rrtconnect_batch_accel(C.byref(desc.as_struct()), accel.h if accel is not None else None,
                               C.c_int(D), C.c_void_p(lo.ctypes.data), C.c_void_p(hi.ctypes.data),
                               C.c_void_p(seg.ctypes.data), C.c_void_p(so2.ctypes.data), C.c_void_p(starts.ctypes.data),
                               C.c_void_p(goals.ctypes.data), C.c_void_p(ng.ctypes.data), C.c_int(G), C.c_int64(Q), C.c_double(range_),
                               C.c_int64(int(max_grow_calls)), C.c_double(time_limit), C.c_uint64(int(seed)), C.c_void_p(status.ctypes.data),
                               C.c_void_p(plen.ctypes.data), C.c_void_p(paths.ctypes.data), C.c_int(max_path), C.byref(gc), C.byref(sc),
                               C.c_int(int(threads)))
    return status, [paths[i, : plen[i]].copy() for i in range(Q)], gc.value, sc.value


def ik_clik_batch(desc, link, poses, q_init, mask=None, eps=1e-5, max_iter=1000, dt=0.1, damp=1e-12, threads=8):
    """computeIKCLIK restatement for n problems -> (q [n, D], success bool[n], err [n, 6])"""
    poses = np.ascontiguousarray(np.asarray(poses, dtype=np.float64).reshape(-1, 7))
    q0 = _states(desc, q_init)
    n, D = q0.shape
    m = None if mask is None or len(mask) == 0 else np.ascontiguousarray(np.asarray(mask, dtype=np.uint8))
    q = np.zeros((n, D))
    ok = np.zeros(n, dtype=np.uint8)
    err = np.zeros((n, 6))
    rc = lib().orc_ik_clik_batch(C.byref(desc.as_struct()), C.c_int(int(link)), C.c_void_p(poses.ctypes.data), C.c_void_p(q0.ctypes.data),
                                 C.c_void_p(m.ctypes.data if m is not None else None), C.c_int64(n), C.c_double(eps), C.c_int(int(max_iter)),
                                 C.c_double(dt), C.c_double(damp), C.c_void_p(q.ctypes.data), C.c_void_p(ok.ctypes.data),
                                 C.c_void_p(err.ctypes.data), C.c_int(int(threads)))
    assert rc == 0
    return q, ok.astype(bool), err


def link_jacobian_batch(desc, link, q, local=False):
    """computeFullJacobian + getLinkJacobian restatement: q [n, D] -> (J [n, 6, D], link placement [n, 3, 4])"""
    q = _states(desc, q)
    n, D = q.shape
    J = np.zeros((n, 6, D))
    T = np.zeros((n, 3, 4))
    rc = lib().orc_link_jacobian_batch(C.byref(desc.as_struct()), C.c_int(int(link)), C.c_void_p(q.ctypes.data), C.c_int64(n),
                                       C.c_int(1 if local else 0), C.c_void_p(J.ctypes.data), C.c_void_p(T.ctypes.data))
    assert rc == 0
    return J, T
