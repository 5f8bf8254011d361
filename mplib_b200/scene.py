"""Flat scene description shared by the C-ABI (include/mpb.h: ``mpb_scene_desc``) and the oracle.

``SceneBuilder`` merges articulations, attached bodies and scene objects of a PlanningWorld
(``src/planning_world.h:138-141``) into the flat tables; ``SceneDesc.as_struct()`` exposes them
as the ctypes mirror of ``mpb_scene_desc`` (pointers into numpy arrays that the SceneDesc keeps
alive).
"""
from __future__ import annotations

import ctypes as C

import numpy as np

GEOM_CONVEX, GEOM_SPHERE, GEOM_BOX, GEOM_CAPSULE, GEOM_CYLINDER, GEOM_MESH, GEOM_OCTREE = range(7)

_PD = C.POINTER(C.c_double)
_PI = C.POINTER(C.c_int32)


class MpbSceneDesc(C.Structure):
    _fields_ = [
        ("n_joints", C.c_int32),
        ("joint_parent", _PI),
        ("joint_placement", _PD),
        ("joint_type", _PI),
        ("joint_axis", _PD),
        ("joint_qidx", _PI),
        ("nq_full", C.c_int32),
        ("q_base", _PD),
        ("state_dim", C.c_int32),
        ("state_map", _PI),
        ("n_links", C.c_int32),
        ("link_joint", _PI),
        ("link_placement", _PD),
        ("n_objects", C.c_int32),
        ("obj_type", _PI),
        ("obj_link", _PI),
        ("obj_local", _PD),
        ("obj_param", _PD),
        ("obj_geom", _PI),
        ("n_hulls", C.c_int32),
        ("hull_vert_off", _PI),
        ("hull_verts", _PD),
        ("hull_face_off", _PI),
        ("hull_faces", _PI),
        ("n_meshes", C.c_int32),
        ("mesh_vert_off", _PI),
        ("mesh_verts", _PD),
        ("mesh_tri_off", _PI),
        ("mesh_tris", _PI),
        ("n_octrees", C.c_int32),
        ("oct_res", _PD),
        ("oct_cell_off", _PI),
        ("oct_cells", _PI),
        ("n_pairs", C.c_int32),
        ("pair_a", _PI),
        ("pair_b", _PI),
    ]


class MpbStats(C.Structure):
    _fields_ = [
        ("n_states", C.c_int64),
        ("pairs_admitted", C.c_int64),
        ("pairs_after_cull", C.c_int64),
        ("pairs_after_spheres", C.c_int64),
        ("narrow_iterations", C.c_int64),
        ("narrow_capped", C.c_int64),
        ("kernel_launches", C.c_int64),
        ("gpu_ms", C.c_float),
        ("gpu_ms_cull", C.c_float),
        ("gpu_ms_narrow", C.c_float),
        ("host_ms_enqueue", C.c_float),
        ("host_ms_total", C.c_float),
        ("gpu_ms_fk", C.c_float),
        ("fk_bytes_per_state", C.c_int32),
        ("async_pending", C.c_int32),
    ]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


def _f64(a, shape):
    a = np.ascontiguousarray(np.asarray(a, dtype=np.float64).reshape(shape))
    return a


def _i32(a, shape=None):
    a = np.asarray(a, dtype=np.int32)
    if shape is not None:
        a = a.reshape(shape)
    return np.ascontiguousarray(a)


def T12(T):
    """4x4 -> row-major 3x4 (12 doubles)."""
    return np.asarray(T, dtype=np.float64)[:3, :4].reshape(12).copy()


class SceneDesc:
    """Immutable bundle of numpy arrays + lazily built ctypes struct."""

    _ARRAYS = {
        "joint_parent": np.int32, "joint_placement": np.float64, "joint_type": np.int32,
        "joint_axis": np.float64, "joint_qidx": np.int32, "q_base": np.float64, "state_map": np.int32,
        "link_joint": np.int32, "link_placement": np.float64,
        "obj_type": np.int32, "obj_link": np.int32, "obj_local": np.float64, "obj_param": np.float64,
        "obj_geom": np.int32,
        "hull_vert_off": np.int32, "hull_verts": np.float64, "hull_face_off": np.int32, "hull_faces": np.int32,
        "mesh_vert_off": np.int32, "mesh_verts": np.float64, "mesh_tri_off": np.int32, "mesh_tris": np.int32,
        "oct_res": np.float64, "oct_cell_off": np.int32, "oct_cells": np.int32,
        "pair_a": np.int32, "pair_b": np.int32,
    }

    def __init__(self, **kw):
        for name, dt in self._ARRAYS.items():
            a = np.ascontiguousarray(np.asarray(kw[name], dtype=dt))
            if a.size == 0:  # keep a valid pointer for empty tables
                a = np.zeros(max(1, a.size), dtype=dt)[:0].copy() if False else a
            setattr(self, name, a)
        self.n_joints = int(self.joint_parent.shape[0])
        self.nq_full = int(self.q_base.shape[0])
        self.state_dim = int(self.state_map.shape[0])
        self.n_links = int(self.link_joint.shape[0])
        self.n_objects = int(self.obj_type.shape[0])
        self.n_hulls = int(self.hull_vert_off.shape[0]) - 1
        self.n_meshes = int(self.mesh_vert_off.shape[0]) - 1
        self.n_octrees = int(self.oct_res.shape[0])
        self.n_pairs = int(self.pair_a.shape[0])
        # names carried for the host layer / tests (not part of the C struct)
        self.link_names = list(kw.get("link_names", []))
        self.object_names = list(kw.get("object_names", []))
        self.pair_kinds = list(kw.get("pair_kinds", []))
        self._struct = None
        self._pad = {}

    def with_pairs(self, pair_a, pair_b, pair_kinds=()):
        kw = {k: getattr(self, k) for k in self._ARRAYS}
        kw.update(pair_a=pair_a, pair_b=pair_b, link_names=self.link_names, object_names=self.object_names,
                  pair_kinds=list(pair_kinds))
        return SceneDesc(**kw)

    def with_q_base(self, q_base):
        kw = {k: getattr(self, k) for k in self._ARRAYS}
        kw.update(q_base=np.asarray(q_base, dtype=np.float64), link_names=self.link_names,
                  object_names=self.object_names, pair_kinds=self.pair_kinds)
        return SceneDesc(**kw)

    def with_objects(self, idx, obj_local=None, obj_param=None):
        """Copy with the pose (row-major 3x4) / parameters of the objects ``idx`` replaced (mpb_scene_update_objects)."""
        kw = {k: getattr(self, k) for k in self._ARRAYS}
        if obj_local is not None:
            loc = self.obj_local.reshape(-1, 12).copy()
            loc[np.asarray(idx, dtype=np.int64)] = np.asarray(obj_local, dtype=np.float64).reshape(-1, 12)
            kw["obj_local"] = loc
        if obj_param is not None:
            par = self.obj_param.reshape(-1, 4).copy()
            par[np.asarray(idx, dtype=np.int64)] = np.asarray(obj_param, dtype=np.float64).reshape(-1, 4)
            kw["obj_param"] = par
        kw.update(link_names=self.link_names, object_names=self.object_names, pair_kinds=self.pair_kinds)
        return SceneDesc(**kw)

    def _ptr(self, name, ctype):
        a = getattr(self, name)
        if a.size == 0:
            a = self._pad.setdefault(name, np.zeros(4, dtype=a.dtype))
        return a.ctypes.data_as(C.POINTER(ctype))

    def as_struct(self) -> MpbSceneDesc:
        if self._struct is None:
            s = MpbSceneDesc()
            for f in ("n_joints", "nq_full", "state_dim", "n_links", "n_objects", "n_hulls", "n_meshes",
                      "n_octrees", "n_pairs"):
                setattr(s, f, getattr(self, f))
            for name, dt in self._ARRAYS.items():
                setattr(s, name, self._ptr(name, C.c_double if dt == np.float64 else C.c_int32))
            self._struct = s
        return self._struct


class SceneBuilder:
    def __init__(self):
        self.jp, self.jpl, self.jt, self.ja, self.jq = [], [], [], [], []
        self.q_base = []
        self.state_map = []
        self.lj, self.lpl, self.link_names = [], [], []
        self.ot, self.ol, self.oloc, self.opar, self.og, self.object_names = [], [], [], [], [], []
        self.hvo, self.hv, self.hfo, self.hf = [0], [], [0], []
        self.mvo, self.mv, self.mto, self.mt = [0], [], [0], []
        self.ores, self.oco, self.oc = [], [0], []
        self._geom_cache = {}
        self.pairs = []
        self.pair_kinds = []

    # -- kinematics ---------------------------------------------------------------------------
    def add_articulation(self, km, q_current, state_q_indices, prefix=""):
        """Append a KinematicModel.  ``q_current``: full qpos in *model* joint order (nv);
        ``state_q_indices``: model-order joint indices written by the batch state, in state order.
        Returns (joint_offset, link_offset)."""
        j0 = len(self.jp)
        l0 = len(self.lj)
        q0 = len(self.q_base)
        for j in range(km.nv):
            par = int(km.joint_parent[j])
            self.jp.append(0 if par == 0 else par + j0)
            self.jpl.append(T12(km.joint_placement[j]))
            self.jt.append(int(km.joint_type[j]))
            self.ja.append(np.asarray(km.joint_axis[j], dtype=np.float64))
            self.jq.append(q0 + j)
        self.q_base.extend(float(v) for v in q_current)
        self.state_map.extend(q0 + int(k) for k in state_q_indices)
        for l in range(len(km.link_names)):
            lj = int(km.link_joint[l])
            self.lj.append(0 if lj == 0 else lj + j0)
            self.lpl.append(T12(km.link_placement[l]))
            self.link_names.append(prefix + km.link_names[l])
        return j0, l0

    # -- geometry pools -----------------------------------------------------------------------
    def add_hull(self, verts, faces, key=None):
        if key is not None and ("hull", key) in self._geom_cache:
            return self._geom_cache[("hull", key)]
        verts = np.asarray(verts, dtype=np.float64).reshape(-1, 3)
        faces = np.asarray(faces, dtype=np.int32).reshape(-1, 3)
        self.hv.append(verts)
        self.hf.append(faces)
        self.hvo.append(self.hvo[-1] + len(verts))
        self.hfo.append(self.hfo[-1] + len(faces))
        idx = len(self.hvo) - 2
        if key is not None:
            self._geom_cache[("hull", key)] = idx
        return idx

    def add_mesh(self, verts, tris, key=None):
        if key is not None and ("mesh", key) in self._geom_cache:
            return self._geom_cache[("mesh", key)]
        verts = np.asarray(verts, dtype=np.float64).reshape(-1, 3)
        tris = np.asarray(tris, dtype=np.int32).reshape(-1, 3)
        self.mv.append(verts)
        self.mt.append(tris)
        self.mvo.append(self.mvo[-1] + len(verts))
        self.mto.append(self.mto[-1] + len(tris))
        idx = len(self.mvo) - 2
        if key is not None:
            self._geom_cache[("mesh", key)] = idx
        return idx

    def add_octree(self, cells, resolution):
        cells = np.asarray(cells, dtype=np.int32).reshape(-1, 3)
        self.oc.append(cells)
        self.oco.append(self.oco[-1] + len(cells))
        self.ores.append(float(resolution))
        return len(self.ores) - 1

    def add_object(self, gtype, link, local_T, param=(0, 0, 0, 0), geom=-1, name=""):
        p = list(param) + [0.0] * (4 - len(param))
        self.ot.append(int(gtype))
        self.ol.append(int(link))
        self.oloc.append(T12(local_T))
        self.opar.append(np.asarray(p[:4], dtype=np.float64))
        self.og.append(int(geom))
        self.object_names.append(name)
        return len(self.ot) - 1

    def add_pair(self, a, b, kind=""):
        self.pairs.append((int(a), int(b)))
        self.pair_kinds.append(kind)

    def finish(self) -> SceneDesc:
        def cat(lst, cols, dt):
            return np.concatenate(lst, axis=0).astype(dt) if lst else np.zeros((0, cols), dtype=dt)

        nj, nl, no = len(self.jp), len(self.lj), len(self.ot)
        pa = [p[0] for p in self.pairs]
        pb = [p[1] for p in self.pairs]
        return SceneDesc(
            joint_parent=_i32(self.jp), joint_placement=_f64(self.jpl, (nj, 12)), joint_type=_i32(self.jt),
            joint_axis=_f64(self.ja, (nj, 3)), joint_qidx=_i32(self.jq), q_base=_f64(self.q_base, (-1,)),
            state_map=_i32(self.state_map), link_joint=_i32(self.lj), link_placement=_f64(self.lpl, (nl, 12)),
            obj_type=_i32(self.ot), obj_link=_i32(self.ol), obj_local=_f64(self.oloc, (no, 12)),
            obj_param=_f64(self.opar, (no, 4)), obj_geom=_i32(self.og),
            hull_vert_off=_i32(self.hvo), hull_verts=cat(self.hv, 3, np.float64),
            hull_face_off=_i32(self.hfo), hull_faces=cat(self.hf, 3, np.int32),
            mesh_vert_off=_i32(self.mvo), mesh_verts=cat(self.mv, 3, np.float64),
            mesh_tri_off=_i32(self.mto), mesh_tris=cat(self.mt, 3, np.int32),
            oct_res=_f64(self.ores, (-1,)), oct_cell_off=_i32(self.oco), oct_cells=cat(self.oc, 3, np.int32),
            pair_a=_i32(pa), pair_b=_i32(pb),
            link_names=self.link_names, object_names=self.object_names, pair_kinds=self.pair_kinds,
        )
