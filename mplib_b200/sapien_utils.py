"""Scene ingestion at the data level: the shape conversion of ``SapienPlanningWorld`` without SAPIEN.

The reference walks ``physx.PhysxRigidBaseComponent.collision_shapes`` (mplib/sapien_utils/conversion.py:326-409) and
turns each PhysX shape into an FCL ``CollisionObject``; articulations are then built with
``ArticulatedModel.create_from_urdf_string(urdf_without_geometry, srdf_string, [(link, [CollisionObject...])])``
(mplib/sapien_utils/conversion.py:83-108, src/articulated_model.cpp:38-67).  SAPIEN and its renderer are unavailable on
this box, so the shapes come in as plain records with the same fields PhysX exposes:

    {"type": "box", "half_size": (hx, hy, hz)}              {"type": "sphere", "radius": r}
    {"type": "capsule", "radius": r, "half_length": h}       {"type": "cylinder", "radius": r, "half_length": h}
    {"type": "convex_mesh", "vertices": [V,3], "triangles": [F,3], "scale": (1,1,1)}
    {"type": "plane"} (skipped with the reference's warning)  {"type": "triangle_mesh"} (NotImplementedError)

each with an optional ``"local_pose": (p[3], q[4] wxyz)``.  The result feeds the same batched device path as every
other world: nothing here computes collisions."""
from __future__ import annotations

import numpy as np

from .pymp import ArticulatedModel, fcl
from .urdf import mat_to_quat, quat_to_mat

# euler2quat(0, pi/2, 0): PhysX capsules / cylinders run along x, FCL's along z (conversion.py:371-374, 382-385)
_X_TO_Z = np.array([np.cos(np.pi / 4), 0.0, np.sin(np.pi / 4), 0.0])


def _pose_T(pose):
    T = np.eye(4)
    if pose is None:
        return T
    p, q = pose
    q = np.asarray(q, dtype=np.float64)
    T[:3, :3] = quat_to_mat(q[0], q[1], q[2], q[3])
    T[:3, 3] = np.asarray(p, dtype=np.float64)
    return T


def convert_col_shapes(shapes, entity_pose=None, is_articulation_link=False, name=""):
    """``SapienPlanningWorld.convert_sapien_col_shape`` on plain shape records -> list of ``fcl.CollisionObject``.

    For an articulation link the objects sit at the shape's local pose (relative to the link); otherwise at
    ``entity_pose * local_pose`` (conversion.py:360-364)."""
    shapes = list(shapes)
    if len(shapes) == 0:
        return []
    if len(shapes) > 1:
        print(f"Got {len(shapes)} shapes for entity '{name}'")
        print("\x1b[33;1m[Warning] only 1 collision shape per component is properly supported.\x1b[0m")
    base = np.eye(4) if is_articulation_link else _pose_T(entity_pose)
    out = []
    for shape in shapes:
        T = base @ _pose_T(shape.get("local_pose"))
        kind = shape["type"]
        if kind == "box":
            geom = fcl.Box(np.asarray(shape["half_size"], dtype=np.float64) * 2)
        elif kind == "capsule":
            geom = fcl.Capsule(float(shape["radius"]), float(shape["half_length"]) * 2)
            T = T @ _pose_T(((0, 0, 0), _X_TO_Z))
        elif kind == "convex_mesh":
            scale = np.asarray(shape.get("scale", (1.0, 1.0, 1.0)), dtype=np.float64)
            assert np.allclose(scale, 1.0), f"Not unit scale {scale}, need to rescale vertices?"
            geom = fcl.Convex(shape["vertices"], shape["triangles"])
        elif kind == "cylinder":
            geom = fcl.Cylinder(float(shape["radius"]), float(shape["half_length"]) * 2)
            T = T @ _pose_T(((0, 0, 0), _X_TO_Z))
        elif kind == "plane":
            print("\x1b[33;1m[Warning] No support for PhysxCollisionShapePlane yet.\x1b[0m")
            continue
        elif kind == "sphere":
            geom = fcl.Sphere(float(shape["radius"]))
        elif kind == "triangle_mesh":
            raise NotImplementedError("Support for TriangleMesh collision is not implemented yet.")
        else:
            raise TypeError(f"Unknown shape type: {kind}")
        out.append(fcl.CollisionObject(geom, T[:3, 3], mat_to_quat(T[:3, :3])))
    return out


def articulation_from_link_shapes(urdf_string, srdf_string, link_shapes, *, joint_names=(), link_names=(), name="", verbose=False):
    """The articulation branch of ``SapienPlanningWorld.__init__`` (conversion.py:83-108): ``link_shapes`` is a list of
    ``(link_name, [shape records])``; links without shapes are left out exactly as the reference does."""
    collision_links = []
    for link_name, shapes in link_shapes:
        objs = convert_col_shapes(shapes, is_articulation_link=True, name=link_name)
        if objs:
            collision_links.append((link_name, objs))
    art = ArticulatedModel.create_from_urdf_string(urdf_string, srdf_string, collision_links, gravity=(0, 0, -9.81),
                                                   joint_names=list(joint_names), link_names=list(link_names), verbose=verbose)
    art.set_name(name)
    return art
