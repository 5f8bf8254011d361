"""``pymp.pinocchio.PinocchioModel``, ``pymp.fcl.FCLModel`` and ``pymp.articulation.ArticulatedModel``
mirrors.  They hold the tables the scene compiler needs and run every numeric query as a batch of
one through libmpb (no host arithmetic path).

Reference: ``src/pinocchio_model.{h,cpp}``, ``src/fcl_model.{h,cpp}``, ``src/articulated_model.{h,cpp}``
and their bindings ``python/pybind_{pinocchio,fcl,articulation}.hpp``.
"""
from __future__ import annotations

import os

import numpy as np

from .. import urdf as U
from ..scene import SceneBuilder
from . import fcl


class PinocchioModel:
    def __init__(self, urdf_filename, gravity=(0, 0, -9.81), verbose=True):
        self._init(U.UrdfTree(U.read_text(urdf_filename)), verbose)

    @classmethod
    def create_from_urdf_string(cls, urdf_string, gravity=(0, 0, -9.81), verbose=True):
        self = cls.__new__(cls)
        self._init(U.UrdfTree(urdf_string), verbose)
        return self

    def _init(self, tree, verbose):
        self.tree = tree
        self.verbose = verbose
        self.km = U.build_kinematic_model(tree)
        self._dev = None
        self._poses = None
        self.set_joint_order(self.get_joint_names(False))
        self.set_link_order(self.get_link_names(False))

    # -- ordering (src/pinocchio_model.cpp:191-238) ---------------------------------------------
    def set_joint_order(self, names):
        names = list(names)
        idx = []
        for n in names:
            if n not in self.km.joint_names:
                raise ValueError(f"{n} is a invalid name in setJointOrder")
            idx.append(self.km.joint_names.index(n))
        if len(idx) != self.km.nv or len(set(idx)) != self.km.nv:
            raise RuntimeError("setJointOrder failed")
        self.user_joint_names = names
        self.joint_user2model = np.asarray(idx, dtype=np.int64)  # 0-based movable joint index
        self.joint_model2user = np.full(self.km.nv, -1, dtype=np.int64)
        self.joint_model2user[self.joint_user2model] = np.arange(len(idx))
        self._dev = None

    def set_link_order(self, names):
        names = list(names)
        idx = []
        for n in names:
            if n not in self.km.link_names:
                raise ValueError(f"{n} is a invalid names in setLinkOrder")
            idx.append(self.km.link_names.index(n))
        self.user_link_names = names
        self.link_user2model = np.asarray(idx, dtype=np.int64)
        self._dev = None

    # -- metadata -------------------------------------------------------------------------------
    def get_joint_names(self, user=True):
        return list(self.user_joint_names) if user else list(self.km.joint_names)

    def get_link_names(self, user=True):
        return list(self.user_link_names) if user else list(self.km.link_names)

    def get_leaf_links(self):
        return list(self.km.leaf_links)

    def _jm(self, index, user):
        return int(self.joint_user2model[index]) if user else int(index) - 1  # model index 0 = universe

    def get_joint_dim(self, index, user=True):
        return 1 if (user or index > 0) else 0

    def get_joint_dims(self, user=True):
        n = self.km.nv
        return np.ones(n, dtype=np.int32) if user else np.array([0] + [1] * n, dtype=np.int32)

    def get_joint_ids(self, user=True):
        n = self.km.nv
        return np.arange(n, dtype=np.int32) if user else np.array([-1] + list(range(n)), dtype=np.int32)

    def get_parents(self, user=True):
        if user:
            return np.array([self.km.joint_parent[j] for j in self.joint_user2model], dtype=np.int32)
        return np.array([0] + [int(p) for p in self.km.joint_parent], dtype=np.int32)

    def get_joint_types(self, user=True):
        if user:
            return [self.km.joint_short[j] for j in self.joint_user2model]
        return ["JointModelRX"] + list(self.km.joint_short)  # universe placeholder like model_.joints[0]

    def get_joint_limits(self, user=True):
        """``src/pinocchio_model.cpp:58-91``: [1,2] per joint; continuous joints report +-pi."""
        idx = list(self.joint_user2model) if user else list(range(self.km.nv))
        out = []
        for j in idx:
            if self.km.joint_type[j] == U.JT_CONTINUOUS:
                out.append(np.array([[-3.14159265359, 3.14159265359]]))
            else:
                out.append(np.array([[self.km.joint_lower[j], self.km.joint_upper[j]]]))
        return out

    def _supports(self, model_joint):
        chain = []
        j = int(model_joint)
        while j != 0:
            chain.append(j)
            j = int(self.km.joint_parent[j - 1])
        return [0] + chain[::-1]

    def get_chain_joint_index(self, end_effector):
        """``src/pinocchio_model.cpp:240-251``"""
        if end_effector not in self.km.link_names:
            raise ValueError(f"unknown link {end_effector}")
        mj = int(self.km.link_joint[self.km.link_names.index(end_effector)])
        return [int(self.joint_model2user[j - 1]) for j in self._supports(mj) if j > 0 and self.joint_model2user[j - 1] != -1]

    def get_chain_joint_name(self, end_effector):
        return [self.user_joint_names[i] for i in self.get_chain_joint_index(end_effector)]

    def get_random_configuration(self):
        lo = np.where(np.isfinite(self.km.joint_lower), self.km.joint_lower, -np.pi)
        hi = np.where(np.isfinite(self.km.joint_upper), self.km.joint_upper, np.pi)
        q_model = lo + (hi - lo) * np.random.random(self.km.nv)
        return q_model[self.joint_user2model]

    def print_frames(self):
        for i, (n, j) in enumerate(zip(self.km.link_names, self.km.link_joint)):
            print(f"Frame {i} {n} {int(j)} BODY")

    # -- FK (batch of one on the device) ---------------------------------------------------------
    def _device_scene(self):
        from .._lib import DeviceScene

        if self._dev is None:
            b = SceneBuilder()
            # state = full qpos in user order
            b.add_articulation(self.km, np.zeros(self.km.nv), [int(j) for j in self.joint_user2model])
            self._dev = DeviceScene(b.finish())
        return self._dev

    def compute_forward_kinematics_batch(self, qpos):
        """[n, nv] user-order qpos -> [n, n_user_links, 7] float32 link poses."""
        q = np.asarray(qpos, dtype=np.float64)
        if q.ndim != 2 or q.shape[1] != self.km.nv:
            raise RuntimeError("Qpos user2pinocchio failed")
        return self._device_scene().fk(q)[:, self.link_user2model, :]

    def compute_forward_kinematics(self, qpos):
        q = np.asarray(qpos, dtype=np.float64).reshape(-1)
        if q.shape[0] != self.km.nv:
            raise RuntimeError("Qpos user2pinocchio failed")
        self._poses = self.compute_forward_kinematics_batch(q[None, :])[0].astype(np.float64)

    def get_link_pose(self, index):
        if not (0 <= index < len(self.user_link_names)):
            raise RuntimeError("The link index is out of bound!")
        if self._poses is None:
            self.compute_forward_kinematics(np.zeros(self.km.nv))
        return self._poses[index].copy()

    def compute_IK_CLIK_batch(self, index, poses, q_init, mask=(), eps=1e-5, maxIter=1000, dt=1e-1, damp=1e-12):
        """n independent ``computeIKCLIK`` problems (``src/pinocchio_model.cpp:377-432``) in one device call:
        poses [n, 7], q_init [n, nv] (user order) -> (qpos [n, nv], success bool[n], error [n, 6])."""
        if not (0 <= index < len(self.user_link_names)):
            raise RuntimeError("The link index is out of bound!")
        return self._device_scene().ik_clik(int(self.link_user2model[index]), poses, q_init, list(mask), eps, maxIter, dt, damp)

    def compute_IK_CLIK(self, index, pose, q_init, mask=(), eps=1e-5, maxIter=1000, dt=1e-1, damp=1e-12):
        """``python/pybind_pinocchio.hpp:54-57``: -> (qpos, success, error)"""
        q, ok, err = self.compute_IK_CLIK_batch(index, np.asarray(pose, dtype=np.float64)[None, :],
                                                np.asarray(q_init, dtype=np.float64)[None, :], mask, eps, maxIter, dt, damp)
        return q[0], bool(ok[0]), err[0]

    def get_link_jacobian_batch(self, index, qpos, local=False):
        """n states at once: qpos [n, nv] (user order) -> J [n, 6, nv], rows (linear, angular), columns in user joint
        order (``getLinkJacobian`` after ``computeFullJacobian``, ``src/pinocchio_model.cpp:335-360``)."""
        if not (0 <= index < len(self.user_link_names)):
            raise RuntimeError("The link index is out of bound!")
        return self._device_scene().link_jacobian(int(self.link_user2model[index]), qpos, local)[0]

    def compute_full_jacobian(self, qpos):
        """``computeFullJacobian``: remembers the state; ``get_link_jacobian`` evaluates the requested link's rows."""
        q = np.asarray(qpos, dtype=np.float64).reshape(-1)
        if q.shape[0] != self.km.nv:
            raise RuntimeError("Qpos user2pinocchio failed")
        self._jac_q = q.copy()

    def get_link_jacobian(self, index, local=False):
        q = getattr(self, "_jac_q", None)
        if q is None:
            raise RuntimeError("compute_full_jacobian has not been called")
        return self.get_link_jacobian_batch(index, q[None, :], local)[0]

    def compute_single_link_local_jacobian(self, qpos, index):
        """``computeSingleLinkLocalJacobian`` (``src/pinocchio_model.cpp:362-375``): the Jacobian in the link's own frame."""
        q = np.asarray(qpos, dtype=np.float64).reshape(-1)
        if q.shape[0] != self.km.nv:
            raise RuntimeError("Qpos user2pinocchio failed")
        return self.get_link_jacobian_batch(index, q[None, :], True)[0]

    def _not_on_path(self, *a, **k):
        raise NotImplementedError("joint-limited CLIK is outside the accelerated path (SURVEY 8f)")

    compute_IK_CLIK_JL = _not_on_path


class FCLModel:
    """``FCLModelTpl``: collision objects per <collision>, pair list, SRDF pruning."""

    def __init__(self, urdf_filename, verbose=True, convex=False):
        self._init(U.UrdfTree(U.read_text(urdf_filename)), U.urdf_dir(urdf_filename), verbose, convex)

    def _init(self, tree, package_dir, verbose, convex):
        self.verbose, self.use_convex = verbose, convex
        self.collision_objects, self.collision_link_names, self.parent_link_names = [], [], []
        self.collision_origin2link = []
        for link_name, parent_name, c in U.collision_dfs(tree):
            if c.gtype == "mesh":
                fn = c.filename
                if convex and ".convex.stl" not in fn:
                    fn = fn + ".convex.stl"
                path = os.path.join(package_dir, fn)
                geom = fcl.load_mesh_as_Convex(path, c.scale) if convex else fcl.load_mesh_as_BVH(path, c.scale)
            elif c.gtype == "cylinder":
                geom = fcl.Cylinder(c.dims[0], c.dims[1])
            elif c.gtype == "box":
                geom = fcl.Box(c.dims)
            elif c.gtype == "sphere":
                geom = fcl.Sphere(c.dims[0])
            else:
                raise ValueError("Unknown geometry type :")
            self.collision_objects.append(fcl.CollisionObject(geom))
            self.collision_link_names.append(link_name)
            self.parent_link_names.append(parent_name)
            self.collision_origin2link.append(c.origin.copy())
        self.set_link_order(self._unique(self.collision_link_names))
        self.collision_pairs = U.self_collision_pairs(self.collision_link_names, self.parent_link_names, True)

    @staticmethod
    def _unique(names):  # std::unique: drops consecutive duplicates only
        out = []
        for n in names:
            if not out or out[-1] != n:
                out.append(n)
        return out

    @classmethod
    def create_from_urdf_string(cls, urdf_string, collision_links, verbose=True):
        """``src/fcl_model.cpp:37-67``: no geometry in the URDF, objects given per link; only the
        same-link exclusion applies to the pair list."""
        self = cls.__new__(cls)
        U.UrdfTree(urdf_string)  # validity check like urdf::parseURDF
        self.verbose, self.use_convex = verbose, False
        self.collision_objects, self.collision_link_names, self.parent_link_names = [], [], []
        self.collision_origin2link = []
        for link_name, objs in collision_links:
            for obj in objs:
                self.collision_objects.append(obj)
                self.collision_link_names.append(link_name)
                self.collision_origin2link.append(obj.T.copy())
        self.set_link_order(self._unique(self.collision_link_names))
        self.collision_pairs = U.self_collision_pairs(self.collision_link_names, None, False)
        return self

    def set_link_order(self, names):
        names = list(names)
        idx = []
        for n in self.collision_link_names:
            if n not in names:
                raise ValueError("The names does not contain link " + n)
            idx.append(names.index(n))
        self.user_link_names = names
        self.collision_link_user_indices = idx

    def get_collision_pairs(self):
        return list(self.collision_pairs)

    def get_collision_objects(self):
        return list(self.collision_objects)

    def get_collision_link_names(self):
        return list(self.collision_link_names)

    def remove_collision_pairs_from_srdf(self, srdf_filename):
        if srdf_filename == "":
            print("No SRDF file provided!")
            return
        if srdf_filename.rsplit(".", 1)[-1] != "srdf":
            raise RuntimeError(srdf_filename + " does not have the right extension.")
        if not os.path.isfile(srdf_filename):
            raise RuntimeError("Cannot open " + srdf_filename)
        self.remove_collision_pairs_from_srdf_string(U.read_text(srdf_filename))

    def remove_collision_pairs_from_srdf_string(self, srdf_string):
        self.collision_pairs = U.remove_pairs_from_srdf(self.collision_pairs, self.collision_link_names, srdf_string)


class ArticulatedModel:
    def __init__(self, urdf_filename, srdf_filename, gravity=(0, 0, -9.81), joint_names=(), link_names=(),
                 verbose=True, convex=False):
        self.pinocchio_model = PinocchioModel(urdf_filename, gravity, verbose)
        self.fcl_model = FCLModel(urdf_filename, verbose, convex)
        self._finish(list(joint_names), list(link_names), verbose)
        self.fcl_model.remove_collision_pairs_from_srdf(srdf_filename)

    @classmethod
    def create_from_urdf_string(cls, urdf_string, srdf_string, collision_links, gravity=(0, 0, -9.81),
                                joint_names=(), link_names=(), verbose=True):
        self = cls.__new__(cls)
        self.pinocchio_model = PinocchioModel.create_from_urdf_string(urdf_string, gravity, verbose)
        self.fcl_model = FCLModel.create_from_urdf_string(urdf_string, collision_links, verbose)
        self._finish(list(joint_names), list(link_names), verbose)
        self.fcl_model.remove_collision_pairs_from_srdf_string(srdf_string)
        return self

    def _finish(self, joint_names, link_names, verbose):
        pm = self.pinocchio_model
        self.verbose = verbose
        self.name = ""
        self.version = 0  # bumped on structure changes (move group, SRDF) so worlds recompile
        self.qpos_version = 0  # bumped by set_qpos so worlds refresh q_base only
        self.user_link_names = link_names if link_names else pm.get_link_names(False)
        self.user_joint_names = joint_names if joint_names else pm.get_joint_names(False)
        pm.set_link_order(self.user_link_names)
        pm.set_joint_order(self.user_joint_names)
        self.fcl_model.set_link_order(self.user_link_names)
        self.current_qpos = np.zeros(pm.km.nv)
        self.set_move_group(self.user_link_names)

    # -- accessors ------------------------------------------------------------------------------
    def get_pinocchio_model(self):
        return self.pinocchio_model

    def get_fcl_model(self):
        return self.fcl_model

    def get_user_link_names(self):
        return list(self.user_link_names)

    def get_user_joint_names(self):
        return list(self.user_joint_names)

    def get_name(self):
        return self.name

    def set_name(self, name):
        self.name = name

    def get_move_group_joint_indices(self):
        return list(self.move_group_user_joints)

    def get_move_group_end_effectors(self):
        return list(self.move_group_end_effectors)

    def get_move_group_joint_names(self):
        return [self.user_joint_names[i] for i in self.move_group_user_joints]

    def set_move_group(self, end_effectors):
        """``src/articulated_model.cpp:76-98``"""
        if isinstance(end_effectors, str):
            end_effectors = [end_effectors]
        self.move_group_end_effectors = list(end_effectors)
        joints = []
        for ee in end_effectors:
            joints = self.pinocchio_model.get_chain_joint_index(ee) + joints
        self.move_group_user_joints = sorted(set(joints))
        self.qpos_dim = len(self.move_group_user_joints)
        self.version += 1

    def get_qpos(self):
        return self.current_qpos.copy()

    def get_qpos_dim(self):
        return self.qpos_dim

    def set_qpos(self, qpos, full=False):
        """``src/articulated_model.cpp:100-127``.  FK itself is deferred to the next device query
        (the device image takes q_base from ``current_qpos``)."""
        q = np.asarray(qpos, dtype=np.float64).reshape(-1)
        if full:
            if q.shape[0] != self.current_qpos.shape[0]:
                raise RuntimeError("Qpos user2pinocchio failed")
            self.current_qpos = q.copy()
        else:
            if q.shape[0] != self.qpos_dim:
                raise RuntimeError(
                    f"Length is not correct, Dim of Q: {self.qpos_dim} ,Len of qpos: {q.shape[0]}")
            self.current_qpos[self.move_group_user_joints] = q
        self.qpos_version += 1

    def update_SRDF(self, srdf):
        self.fcl_model.remove_collision_pairs_from_srdf(srdf)
        self.version += 1

    # -- scene-compiler helpers -----------------------------------------------------------------
    def q_model_order(self):
        q = np.zeros(self.pinocchio_model.km.nv)
        q[self.pinocchio_model.joint_user2model] = self.current_qpos
        return q

    def state_model_joints(self):
        return [int(self.pinocchio_model.joint_user2model[u]) for u in self.move_group_user_joints]
