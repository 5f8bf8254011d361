"""``pymp.planning_world.PlanningWorld`` mirror (``src/planning_world.{h,cpp}``,
``python/pybind_planning_world.hpp:31-147``).

The world is a host-side container; every collision / distance query compiles (and caches) a flat
scene image and runs on the GPU through libmpb.  Legacy single-state calls (``collide``,
``collide_full`` ...) keep the stateful semantics by running a batch of one against the articulations'
current qpos; the additive ``*_batch`` methods take the states explicitly.

Canonical result order (the reference's is unordered_map-dependent): the reference's loop
order with insertion order standing in for unordered_map iteration.
"""
from __future__ import annotations

import numpy as np

from ..scene import GEOM_CONVEX, GEOM_MESH, GEOM_OCTREE, SceneBuilder
from ..urdf import posevec_to_T
from . import fcl
from .collision_matrix import AllowedCollisionMatrix


class WorldCollisionResult:
    def __init__(self, res=None, collision_type="", object_name1="", object_name2="", link_name1="", link_name2=""):
        self.res = res if res is not None else fcl.CollisionResult()
        self.collision_type = collision_type
        self.object_name1, self.object_name2 = object_name1, object_name2
        self.link_name1, self.link_name2 = link_name1, link_name2

    def key(self):
        return (self.collision_type, self.object_name1, self.object_name2, self.link_name1, self.link_name2)

    def __repr__(self):
        return f"WorldCollisionResult{self.key()}"


class WorldDistanceResult:
    def __init__(self):
        self.res = fcl.DistanceResult()
        self.min_distance = float(np.finfo(np.float64).max)
        self.distance_type = ""
        self.object_name1 = self.object_name2 = self.link_name1 = self.link_name2 = ""


class AttachedBody:
    """``src/attached_body.{h,cpp}``: object rigidly attached to ``link_id`` (user link index)."""

    def __init__(self, name, obj, attached_articulation, attached_link_id, pose, touch_links=()):
        self.name, self.object = name, obj
        self.attached_articulation = attached_articulation
        self.attached_link_id = int(attached_link_id)
        self.pose = np.asarray(pose, dtype=np.float64)  # 4x4
        self.touch_links = list(touch_links)

    def get_name(self):
        return self.name

    def get_object(self):
        return self.object

    def get_attached_articulation(self):
        return self.attached_articulation

    def get_attached_link_id(self):
        return self.attached_link_id

    def get_pose(self):
        return self.pose.copy()

    def set_pose(self, pose):
        self.pose = np.asarray(pose, dtype=np.float64)

    def get_touch_links(self):
        return list(self.touch_links)

    def set_touch_links(self, touch_links):
        self.touch_links = list(touch_links)


class PlanningWorld:
    def __init__(self, articulations, articulation_names, normal_objects=(), normal_object_names=(), device=0):
        if len(articulations) != len(articulation_names):
            raise RuntimeError("articulations and articulation_names should have the same size")
        if len(normal_objects) != len(normal_object_names):
            raise RuntimeError("normal_objects and normal_object_names should have the same size")
        self.acm = AllowedCollisionMatrix()
        self.articulations, self.planned = {}, set()
        self.normal_objects, self.attached_bodies = {}, {}
        self.device = device
        for a, n in zip(articulations, articulation_names):
            a.set_name(n)
            self.articulations[n] = a
            self.planned.add(n)
        for o, n in zip(normal_objects, normal_object_names):
            self.normal_objects[n] = o
        self._version = 0
        self._cache = None

    # -- articulations --------------------------------------------------------------------------
    def get_articulation_names(self):
        return list(self.articulations.keys())

    def get_planned_articulations(self):
        return [self.articulations[n] for n in sorted(self.planned)]

    def get_articulation(self, name):
        return self.articulations.get(name)

    def has_articulation(self, name):
        return name in self.articulations

    def add_articulation(self, name, model, planned=False):
        model.set_name(name)
        self.articulations[name] = model
        self.set_articulation_planned(name, planned)
        self._version += 1

    def remove_articulation(self, name):
        art = self.articulations.pop(name, None)
        if art is None:
            return False
        self.planned.discard(name)
        self.acm.remove_entry(art.get_user_link_names())
        self.acm.remove_default_entry(art.get_user_link_names())
        self._version += 1
        return True

    def is_articulation_planned(self, name):
        return name in self.planned

    def set_articulation_planned(self, name, planned):
        if name not in self.articulations:
            raise IndexError(name)  # .at() -> std::out_of_range
        if planned:
            self.planned.add(name)
        else:
            self.planned.discard(name)
        self._version += 1

    # -- objects --------------------------------------------------------------------------------
    def get_normal_object_names(self):
        return list(self.normal_objects.keys())

    def get_normal_object(self, name):
        return self.normal_objects.get(name)

    def has_normal_object(self, name):
        return name in self.normal_objects

    def add_normal_object(self, name, collision_object):
        self.normal_objects[name] = collision_object
        self._version += 1

    def add_point_cloud(self, name, vertices, resolution=0.01):
        """``src/planning_world.cpp:101-110``"""
        self.add_normal_object(name, fcl.CollisionObject(fcl.OcTree(vertices, resolution)))

    def remove_normal_object(self, name):
        if name not in self.normal_objects:
            return False
        del self.normal_objects[name]
        self.attached_bodies.pop(name, None)
        self.acm.remove_entry(name)
        self.acm.remove_default_entry(name)
        self._version += 1
        return True

    # -- attach / detach (src/planning_world.cpp:123-234) -------------------------------------------
    def is_normal_object_attached(self, name):
        return name in self.attached_bodies

    def get_attached_object(self, name):
        return self.attached_bodies.get(name)

    def attach_object(self, name, *args):
        """Overloads: (name, art_name, link_id, pose[, touch_links]) and
        (name, p_geom, art_name, link_id, pose[, touch_links])."""
        if args and isinstance(args[0], fcl.CollisionGeometry):
            geom, args = args[0], args[1:]
            self.remove_normal_object(name)
            self.add_normal_object(name, fcl.CollisionObject(geom))
        art_name, link_id, pose = args[0], int(args[1]), np.asarray(args[2], dtype=np.float64).reshape(-1)
        touch_links = list(args[3]) if len(args) > 3 else None
        if name not in self.normal_objects:
            raise IndexError(name)
        if art_name not in self.planned:
            raise IndexError(art_name)
        obj, art = self.normal_objects[name], self.articulations[art_name]
        prev = self.attached_bodies.pop(name, None)
        if touch_links is not None:
            body = AttachedBody(name, obj, art, link_id, posevec_to_T(pose), touch_links)
            if prev is not None:
                self.acm.remove_entry(name, prev.get_touch_links())
            self.attached_bodies[name] = body
            self.acm.set_entry(name, touch_links, True)
        else:
            body = AttachedBody(name, obj, art, link_id, posevec_to_T(pose))
            self.attached_bodies[name] = body
            if prev is not None:
                body.set_touch_links(prev.get_touch_links())
            else:
                self._version += 1
                tl = []
                for c in self.self_collide():
                    if c.link_name1 == name:
                        tl.append(c.link_name2)
                    elif c.link_name2 == name:
                        tl.append(c.link_name1)
                body.set_touch_links(tl)
                self.acm.set_entry(name, tl, True)
        self._version += 1

    def attach_sphere(self, radius, art_name, link_id, pose):
        self.attach_object(f"{art_name}_{link_id}_sphere", fcl.Sphere(radius), art_name, link_id, pose)

    def attach_box(self, size, art_name, link_id, pose):
        self.attach_object(f"{art_name}_{link_id}_box", fcl.Box(size), art_name, link_id, pose)

    def attach_mesh(self, mesh_path, art_name, link_id, pose):
        self.attach_object(f"{art_name}_{link_id}_mesh", fcl.load_mesh_as_BVH(mesh_path, (1, 1, 1)), art_name, link_id, pose)

    def detach_object(self, name, also_remove=False):
        if also_remove:
            self.normal_objects.pop(name, None)
            self.acm.remove_entry(name)
            self.acm.remove_default_entry(name)
        self._version += 1
        body = self.attached_bodies.pop(name, None)
        if body is None:
            return False
        self.acm.remove_entry(name, body.get_touch_links())
        return True

    def print_attached_body_pose(self):
        for name, body in self.attached_bodies.items():
            print(name, "attached to link", body.attached_link_id, "pose:\n", body.pose)

    # -- state ----------------------------------------------------------------------------------
    def set_qpos(self, name, qpos):
        if name not in self.articulations:
            raise IndexError(name)
        self.articulations[name].set_qpos(qpos)

    def set_qpos_all(self, state):
        """``src/planning_world.cpp:249-262``: planned articulations in name order."""
        state = np.asarray(state, dtype=np.float64).reshape(-1)
        i = 0
        for n in sorted(self.planned):
            art = self.articulations[n]
            d = art.get_qpos_dim()
            seg = state[i: i + d]
            if seg.shape[0] != d:
                raise RuntimeError(f"Bug with size {seg.shape[0]} {d}")
            art.set_qpos(seg)
            i += d
        if i != state.shape[0]:
            raise RuntimeError("State dimension is not correct")

    def get_allowed_collision_matrix(self):
        return self.acm

    # -- scene compiler -------------------------------------------------------------------------
    def _geom_of(self, b, geom):
        """-> (gtype, param, geom index) registering pooled geometry in the builder."""
        gt = geom.gtype
        if gt == GEOM_CONVEX:
            return gt, (0, 0, 0, 0), b.add_hull(geom.vertices, geom.faces, key=getattr(geom, "_key", None) or id(geom))
        if gt == GEOM_MESH:
            return gt, (0, 0, 0, 0), b.add_mesh(geom.vertices, geom.tris, key=getattr(geom, "_key", None) or id(geom))
        if gt == GEOM_OCTREE:
            return gt, (0, 0, 0, 0), b.add_octree(geom.cells, geom.resolution)
        if gt < 0:
            raise NotImplementedError(f"{type(geom).__name__} is outside the accelerated geometry set")
        return gt, geom._param(), -1

    def _structure_sig(self):
        """What forces a re-compile: the set of articulations / objects / attachments and their geometry.  Poses are not
        part of it (``_pose_sig``): a moved obstacle or a re-posed attached body is an in-place update."""
        arts = tuple((n, id(a), a.version, n in self.planned) for n, a in sorted(self.articulations.items()))
        objs = tuple((n, id(o), id(o.geometry), self._geom_sig(o.geometry)) for n, o in self.normal_objects.items())
        att = tuple((n, id(bd.attached_articulation), bd.attached_link_id) for n, bd in self.attached_bodies.items())
        return (self._version, arts, objs, att)

    @staticmethod
    def _geom_sig(g):
        return tuple(np.asarray(g._param(), dtype=np.float64).tolist()) if 0 < g.gtype < GEOM_MESH else ()

    def _pose_sig(self):
        return tuple((n, (self.attached_bodies[n].pose if n in self.attached_bodies else o.T).tobytes())
                     for n, o in self.normal_objects.items())

    def compile(self):
        """Flatten the world (see SceneBuilder).  Returns (SceneDesc, candidate pairs with metadata)."""
        b = SceneBuilder()
        art_objs = {}  # art name -> list of object indices (FCL collision object order)
        art_link0 = {}
        for name in sorted(self.articulations):
            art = self.articulations[name]
            pm, fm = art.get_pinocchio_model(), art.get_fcl_model()
            state_joints = art.state_model_joints() if name in self.planned else []
            _, l0 = b.add_articulation(pm.km, art.q_model_order(), state_joints, prefix="")
            art_link0[name] = l0
            idxs = []
            for i, obj in enumerate(fm.collision_objects):
                gt, prm, gi = self._geom_of(b, obj.geometry)
                link = l0 + int(pm.link_user2model[fm.collision_link_user_indices[i]])
                idxs.append(b.add_object(gt, link, fm.collision_origin2link[i], prm, gi, name=fm.collision_link_names[i]))
            art_objs[name] = idxs
        att_obj, scene_obj = {}, {}
        for oname, obj in self.normal_objects.items():
            gt, prm, gi = self._geom_of(b, obj.geometry)
            if oname in self.attached_bodies:
                bd = self.attached_bodies[oname]
                art = bd.attached_articulation
                pm = art.get_pinocchio_model()
                link = art_link0[art.get_name()] + int(pm.link_user2model[bd.attached_link_id])
                att_obj[oname] = b.add_object(gt, link, bd.pose, prm, gi, name=oname)
            else:
                scene_obj[oname] = b.add_object(gt, -1, obj.T, prm, gi, name=oname)

        # candidate pairs in the reference's loop order, with the result metadata each would carry
        cand = []  # (a, b, section, type, object_name1, object_name2, link_name1, link_name2)
        planned = sorted(self.planned)
        unplanned = [n for n in self.articulations if n not in self.planned]
        for an in planned:  # selfCollide, src/planning_world.cpp:284-348
            fm = self.articulations[an].get_fcl_model()
            ln, ob = fm.collision_link_names, art_objs[an]
            for (x, y) in fm.collision_pairs:
                cand.append((ob[x], ob[y], "self", "self", an, an, ln[x], ln[y]))
            for an2 in planned:
                if an2 == an:
                    break
                fm2 = self.articulations[an2].get_fcl_model()
                for i in range(len(ob)):
                    for j in range(len(art_objs[an2])):
                        cand.append((ob[i], art_objs[an2][j], "self", "self_articulation", an, an2, ln[i],
                                     fm2.collision_link_names[j]))
            for bn, bo in att_obj.items():
                for i in range(len(ob)):
                    cand.append((bo, ob[i], "self", "self_attach", an, bn, ln[i], bn))
        names = list(att_obj.keys())  # attach_attach, :350-367
        for i1 in range(len(names)):
            for i2 in range(i1):
                n1, n2 = names[i1], names[i2]
                cand.append((att_obj[n1], att_obj[n2], "self", "attach_attach", n1, n2, n1, n2))
        for an in planned:  # collideWithOthers, :389-435
            fm = self.articulations[an].get_fcl_model()
            ln, ob = fm.collision_link_names, art_objs[an]
            for an2 in unplanned:
                fm2 = self.articulations[an2].get_fcl_model()
                for i in range(len(ob)):
                    for j in range(len(art_objs[an2])):
                        cand.append((ob[i], art_objs[an2][j], "others", "articulation_articulation", an, an2, ln[i],
                                     fm2.collision_link_names[j]))
            for sn, so in scene_obj.items():
                for i in range(len(ob)):
                    cand.append((ob[i], so, "others", "articulation_sceneobject", an, sn, ln[i], sn))
        for bn, bo in att_obj.items():  # :437-479
            for an2 in unplanned:
                fm2 = self.articulations[an2].get_fcl_model()
                for j in range(len(art_objs[an2])):
                    cand.append((bo, art_objs[an2][j], "others", "attach_articulation", bn, an2, bn,
                                 fm2.collision_link_names[j]))
            for sn, so in scene_obj.items():
                cand.append((bo, so, "others", "attach_sceneobject", bn, sn, bn, sn))
        self._obj_index = dict(scene_obj)
        self._obj_index.update(att_obj)
        return b, cand

    def _admitted(self, cand):
        """ACM hoisted in front of the narrowphase (equals filterCollisions, src/planning_world.cpp:264-274)."""
        return [c for c in cand if self.acm.admits(c[6], c[7])]

    def _sync(self):
        from .._lib import DeviceScene

        sig = self._structure_sig()
        c = self._cache
        if c is None or c["sig"] != sig:
            if c is not None:
                c["dev"].close()
            b, cand = self.compile()
            pairs = self._admitted(cand)
            for p in pairs:
                b.add_pair(p[0], p[1], p[3])
            desc = b.finish()
            self._cache = c = {"sig": sig, "dev": DeviceScene(desc, self.device), "cand": cand, "pairs": pairs,
                               "acm": self.acm.version, "q": self._q_sig(), "pose": self._pose_sig(), "obj_index": self._obj_index}
        elif c["acm"] != self.acm.version:
            pairs = self._admitted(c["cand"])
            c["dev"].set_pairs([p[0] for p in pairs], [p[1] for p in pairs], [p[3] for p in pairs])
            c["pairs"], c["acm"] = pairs, self.acm.version
        ps = self._pose_sig()
        if c["pose"] != ps:  # CollisionObject.set_transformation / a re-posed attached body: in-place update, no re-compile
            old = dict(c["pose"])
            idx, loc = [], []
            for n, t in ps:
                if old.get(n) != t:
                    T = self.attached_bodies[n].pose if n in self.attached_bodies else self.normal_objects[n].T
                    idx.append(c["obj_index"][n])
                    loc.append(np.asarray(T, dtype=np.float64)[:3, :4].reshape(12))
            if idx:
                c["dev"].update_objects(idx, np.stack(loc))
            c["pose"] = ps
        if c["q"] != self._q_sig():
            qb = [self.articulations[n].q_model_order() for n in sorted(self.articulations)]
            c["dev"].set_q_base(np.concatenate(qb) if qb else np.zeros(0))
            c["q"] = self._q_sig()
        return c

    def _q_sig(self):
        return tuple((n, a.qpos_version) for n, a in sorted(self.articulations.items()))

    def get_scene_desc(self):
        """The compiled flat description (used by tests and the oracle)."""
        return self._sync()["dev"].desc

    def get_pair_metadata(self):
        return [p[2:] for p in self._sync()["pairs"]]

    def _current_state(self):
        parts = [self.articulations[n].get_qpos()[self.articulations[n].get_move_group_joint_indices()] for n in sorted(self.planned)]
        return np.concatenate(parts) if parts else np.zeros(0)

    # -- batch API (additive) ---------------------------------------------------------------------
    def collide_full_batch(self, states, want_pairs=True, stats=False):
        """states [n, D] -> (verdict bool[n], pair_hit bool[n, P]) over the admitted pair list
        (``get_pair_metadata()`` gives (section, type, names) per column)."""
        c = self._sync()
        states = np.asarray(states)
        if states.ndim != 2 or states.shape[1] != c["dev"].state_dim:
            raise RuntimeError("State dimension is not correct")
        return c["dev"].collide(states, want_pairs=want_pairs, stats=stats)

    def collide_batch(self, states):
        """Batched ``ValidityCheckerTpl::isValid`` complement: True where the state collides."""
        return self.collide_full_batch(states, want_pairs=False)[0]

    # -- legacy single-state queries ----------------------------------------------------------------
    def _query(self, section):
        c = self._sync()
        _, hit = c["dev"].collide(self._current_state()[None, :], want_pairs=True)
        out = []
        for k, p in enumerate(c["pairs"]):
            if hit[0, k] and (section is None or p[2] == section):
                out.append(WorldCollisionResult(fcl.CollisionResult(True), p[3], p[4], p[5], p[6], p[7]))
        return out

    def self_collide(self, request=None):
        return self._query("self")

    def collide_with_others(self, request=None):
        return self._query("others")

    def collide_full(self, request=None):
        r = self._query(None)
        return [x for x in r if x.collision_type in ("self", "self_articulation", "self_attach", "attach_attach")] + \
               [x for x in r if x.collision_type not in ("self", "self_articulation", "self_attach", "attach_attach")]

    def collide(self, request=None):
        return len(self.collide_full(request)) > 0

    # -- distance (src/planning_world.cpp:492-720) ---------------------------------------------------------
    def distance_batch(self, states, stats=False):
        """Batched ``ValidityCheckerTpl::clearance``: states [n, D] -> (min_distance float64[n], argmin pair int32[n]);
        ``get_pair_metadata()[argmin]`` names the pair.  -1 where an admitted pair touches / intersects."""
        c = self._sync()
        states = np.asarray(states)
        if states.ndim != 2 or states.shape[1] != c["dev"].state_dim:
            raise RuntimeError("State dimension is not correct")
        return c["dev"].distance(states, stats=stats)

    def _distance_query(self, section):
        c = self._sync()
        pairs = c["pairs"]
        sel = [k for k, p in enumerate(pairs) if section is None or p[2] == section]
        dev = c["dev"]
        if len(sel) != len(pairs):  # query a sub-list of the admitted pairs, then restore the full list
            dev.set_pairs([pairs[k][0] for k in sel], [pairs[k][1] for k in sel], [pairs[k][3] for k in sel])
        try:
            md, am = dev.distance(self._current_state()[None, :])
        finally:
            if len(sel) != len(pairs):
                dev.set_pairs([p[0] for p in pairs], [p[1] for p in pairs], [p[3] for p in pairs])
        r = WorldDistanceResult()
        if am[0] >= 0:
            p = pairs[sel[int(am[0])]]
            r.min_distance = float(md[0])
            r.res = fcl.DistanceResult(float(md[0]))
            r.distance_type, r.object_name1, r.object_name2, r.link_name1, r.link_name2 = p[3], p[4], p[5], p[6], p[7]
        return r

    def self_distance(self, request=None):
        return self._distance_query("self")

    def distance_with_others(self, request=None):
        return self._distance_query("others")

    def distance_full(self, request=None):
        return self._distance_query(None)

    def distance(self, request=None):
        return self.distance_full(request).min_distance
