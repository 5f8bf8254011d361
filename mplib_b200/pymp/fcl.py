"""``pymp.fcl`` mirror: collision geometry, objects, requests and results.

Host-side data holders with the names / constructor arguments of the reference bindings
(``python/pybind_fcl.hpp:59-436``).  No arithmetic lives here: queries go through the C-ABI
(libmpb.so) via the scene compiler.  Geometry classes outside the hot path's type set (Cone,
Ellipsoid, Halfspace, Plane, TriangleP) are declared so construction works; putting one into a
world raises ``NotImplementedError`` from the scene compiler.
"""
from __future__ import annotations

import numpy as np

from .. import stl as _stl
from ..urdf import quat_to_mat, posevec_to_T, mat_to_quat
from ..scene import (GEOM_BOX, GEOM_CAPSULE, GEOM_CONVEX, GEOM_CYLINDER, GEOM_MESH, GEOM_OCTREE, GEOM_SPHERE)


def _vec(v, n):
    a = np.asarray(v, dtype=np.float64).reshape(-1)
    if a.shape[0] != n:
        raise TypeError(f"expected a vector of length {n}")
    return a


class Triangle:
    def __init__(self, p1=0, p2=0, p3=0):
        self._v = [int(p1), int(p2), int(p3)]

    def set(self, p1, p2, p3):
        self._v = [int(p1), int(p2), int(p3)]

    def get(self, i):
        return self._v[i]

    __getitem__ = get


class CollisionGeometry:
    gtype = -1
    cost_density = 1.0

    def _param(self):
        return (0.0, 0.0, 0.0, 0.0)


class Box(CollisionGeometry):
    gtype = GEOM_BOX

    def __init__(self, *args, **kw):
        if "side" in kw:
            args = (kw["side"],)
        elif kw:
            args = (kw["x"], kw["y"], kw["z"])
        self.side = _vec(args[0], 3) if len(args) == 1 else _vec(args, 3)

    def _param(self):
        return (self.side[0], self.side[1], self.side[2], 0.0)


class Sphere(CollisionGeometry):
    gtype = GEOM_SPHERE

    def __init__(self, radius):
        self.radius = float(radius)

    def _param(self):
        return (self.radius, 0.0, 0.0, 0.0)


class Capsule(CollisionGeometry):
    gtype = GEOM_CAPSULE

    def __init__(self, radius, lz):
        self.radius, self.lz = float(radius), float(lz)

    def _param(self):
        return (self.radius, self.lz, 0.0, 0.0)


class Cylinder(CollisionGeometry):
    gtype = GEOM_CYLINDER

    def __init__(self, radius, lz):
        self.radius, self.lz = float(radius), float(lz)

    def _param(self):
        return (self.radius, self.lz, 0.0, 0.0)


class _Unsupported(CollisionGeometry):
    def __init__(self, *a, **k):
        self.args = (a, k)


class Cone(_Unsupported):
    pass


class Ellipsoid(_Unsupported):
    pass


class Halfspace(_Unsupported):
    pass


class Plane(_Unsupported):
    pass


class TriangleP(_Unsupported):
    pass


class Convex(CollisionGeometry):
    """``fcl::Convex`` from (vertices[V,3], faces[F,3]) (``python/pybind_fcl.hpp:104-124``)."""

    gtype = GEOM_CONVEX

    def __init__(self, vertices, faces, throw_if_invalid=True):
        self.vertices = np.ascontiguousarray(np.asarray(vertices, dtype=np.float64).reshape(-1, 3))
        self.faces = np.ascontiguousarray(np.asarray(faces, dtype=np.int32).reshape(-1, 3))
        if self.vertices.shape[0] < 4 and throw_if_invalid:
            raise ValueError("Convex needs at least 4 vertices")
        if self.faces.size and (self.faces.min() < 0 or self.faces.max() >= self.vertices.shape[0]):
            raise ValueError("Convex face index out of range")
        self._key = None

    def get_face_count(self):
        return int(self.faces.shape[0])

    def get_faces(self):
        out = []
        for f in self.faces:
            out.extend([3, int(f[0]), int(f[1]), int(f[2])])
        return out

    def get_vertices(self):
        return [v.copy() for v in self.vertices]

    def get_interior_point(self):
        return self.vertices.mean(axis=0)

    def compute_volume(self):
        v = self.vertices
        f = self.faces
        return float(abs(np.einsum("ij,ij->i", v[f[:, 0]], np.cross(v[f[:, 1]], v[f[:, 2]])).sum()) / 6.0)


class BVHModel(CollisionGeometry):
    """``fcl::BVHModel<OBBRSS>`` built with beginModel/addSubModel/endModel."""

    gtype = GEOM_MESH

    def __init__(self):
        self.vertices = np.zeros((0, 3), dtype=np.float64)
        self.tris = np.zeros((0, 3), dtype=np.int32)
        self._key = None

    def beginModel(self, num_faces=0, num_vertices=0):
        return 0

    def endModel(self):
        return 0

    def addSubModel(self, vertices, faces=None):
        v = np.asarray([np.asarray(x, dtype=np.float64) for x in vertices], dtype=np.float64).reshape(-1, 3)
        off = self.vertices.shape[0]
        self.vertices = np.concatenate([self.vertices, v], axis=0)
        if faces is not None:
            f = np.asarray([[t[0], t[1], t[2]] for t in faces], dtype=np.int32).reshape(-1, 3) + off
            self.tris = np.concatenate([self.tris, f], axis=0)
        return 0

    def get_vertices(self):
        return [v.copy() for v in self.vertices]

    def get_faces(self):
        return [Triangle(*t) for t in self.tris]

    @property
    def num_faces(self):
        return int(self.tris.shape[0])

    @property
    def num_vertices(self):
        return int(self.vertices.shape[0])


def octomap_cells(vertices, resolution):
    """Occupied finest-level voxels of ``octomap::OcTree(res)`` after ``updateNode(p, true)`` per
    point (``src/planning_world.cpp:101-110``).  [upstream octomap 1.9.8] ``coordToKey`` is
    ``floor(coord / res) + 32768``; one hit gives log-odds 0.85 > 0 so every touched voxel is
    occupied; pruning merges 8 equal children into one box covering the same volume."""
    v = np.asarray(vertices, dtype=np.float64).reshape(-1, 3)
    # octomap::point3d stores float32 coordinates
    v = v.astype(np.float32).astype(np.float64)
    cells = np.floor(v / float(resolution)).astype(np.int64)
    if cells.size and (np.abs(cells).max() >= 32768):
        cells = cells[(np.abs(cells) < 32768).all(axis=1)]  # updateNode ignores out-of-range keys
    cells = np.unique(cells, axis=0) if cells.size else cells.reshape(0, 3)
    return cells.astype(np.int32)


class OcTree(CollisionGeometry):
    gtype = GEOM_OCTREE

    def __init__(self, vertices=None, resolution=None):
        if resolution is None:  # OcTree(resolution)
            resolution, vertices = vertices, None
        self.resolution = float(resolution)
        self.cells = octomap_cells(vertices, self.resolution) if vertices is not None else np.zeros((0, 3), np.int32)


def load_mesh_as_BVH(mesh_path, scale=(1.0, 1.0, 1.0)):
    """``src/urdf_utils.cpp:135-154``"""
    v, t = _stl.load_stl(mesh_path, scale)
    m = BVHModel()
    m.vertices, m.tris = v, t
    m._key = ("file", mesh_path, tuple(float(s) for s in scale))
    return m


def load_mesh_as_Convex(mesh_path, scale=(1.0, 1.0, 1.0)):
    """``src/urdf_utils.cpp:156-183``"""
    v, t = _stl.load_stl(mesh_path, scale)
    c = Convex(v, t, True)
    c._key = ("file", mesh_path, tuple(float(s) for s in scale))
    return c


class CollisionObject:
    """Geometry + pose (``python/pybind_fcl.hpp:239-259``); pose is mutable after insertion into a
    world, exactly like the shared_ptr the reference world keeps."""

    def __init__(self, collision_geometry, position=(0, 0, 0), quaternion=(1, 0, 0, 0)):
        self.geometry = collision_geometry
        q = _vec(quaternion, 4)
        self.T = np.eye(4)
        self.T[:3, :3] = quat_to_mat(q[0], q[1], q[2], q[3])
        self.T[:3, 3] = _vec(position, 3)

    def get_collision_geometry(self):
        return self.geometry

    def get_translation(self):
        return self.T[:3, 3].copy()

    def get_rotation(self):
        return self.T[:3, :3].copy()

    def set_transformation(self, pose):
        self.T = posevec_to_T(_vec(pose, 7))

    def get_pose7(self):
        q = mat_to_quat(self.T[:3, :3])
        return np.concatenate([self.T[:3, 3], q])


GST_LIBCCD, GST_INDEP = 0, 1


class GJKSolverType:
    GST_LIBCCD = GST_LIBCCD
    GST_INDEP = GST_INDEP


class CollisionRequest:
    def __init__(self, num_max_contacts=1, enable_contact=False, num_max_cost_sources=1, enable_cost=False,
                 use_approximate_cost=True, gjk_solver_type=GST_LIBCCD, gjk_tolerance=1e-6):
        self.num_max_contacts = num_max_contacts
        self.enable_contact = enable_contact
        self.num_max_cost_sources = num_max_cost_sources
        self.enable_cost = enable_cost
        self.use_approximate_cost = use_approximate_cost
        self.gjk_solver_type = gjk_solver_type
        self.gjk_tolerance = gjk_tolerance

    def isSatisfied(self, result):
        return result.num_contacts() >= self.num_max_contacts


class Contact:
    def __init__(self):
        self.penetration_depth = 0.0
        self.normal = np.zeros(3)
        self.pos = np.zeros(3)


class CollisionResult:
    """Boolean outcome of one pair; contacts are not produced (the default request has
    ``enable_contact=False``, ``python/pybind_fcl.hpp:272-277``)."""

    def __init__(self, colliding=False):
        self._hit = bool(colliding)

    def is_collision(self):
        return self._hit

    def num_contacts(self):
        return 1 if self._hit else 0

    def num_cost_sources(self):
        return 0

    def get_contacts(self):
        return [Contact()] if self._hit else []

    def get_contact(self, i):
        return self.get_contacts()[i]

    def get_cost_sources(self):
        return []

    def clear(self):
        self._hit = False


class DistanceRequest:
    def __init__(self, enable_nearest_points=False, enable_signed_distance=False, rel_err=0.0, abs_err=0.0,
                 distance_tolerance=1e-6, gjk_solver_type=GST_LIBCCD):
        self.enable_nearest_points = enable_nearest_points
        self.enable_signed_distance = enable_signed_distance
        self.rel_err, self.abs_err = rel_err, abs_err
        self.distance_tolerance = distance_tolerance
        self.gjk_solver_type = gjk_solver_type

    def isSatisfied(self, result):
        return False


class DistanceResult:
    def __init__(self, min_distance=float("inf")):
        self.min_distance = float(min_distance)
        self.nearest_points = [np.zeros(3), np.zeros(3)]

    def clear(self):
        self.min_distance = float("inf")
