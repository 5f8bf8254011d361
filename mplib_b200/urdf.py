"""URDF / SRDF -> flat kinematic and collision tables (host side of the scene compiler).

Restates, without urdfdom / Pinocchio / Boost, the model-building conventions of the
reference so that joint order, link order, collision-object order and the self-collision
pair list come out identical:

* URDF tree + child ordering: urdfdom keeps joints in a ``std::map`` keyed by joint name, so
  a link's children are visited in joint-name order.  DFS as in
  ``src/pinocchio_model.cpp:558-726`` and ``src/fcl_model.cpp:195-265``.
* Pinocchio-style model (``src/pinocchio_model.cpp:728-752``): only non-fixed joints become
  model joints; a fixed joint turns into a frame on the last movable ancestor with the
  accumulated placement.  ``oMi[j] = oMi[parent[j]] * placement[j] * J_j(q_j)``.
* rpy -> rotation follows urdfdom (``Rotation::setFromRPY``: quaternion from half angles),
  then quaternion -> matrix as ``src/urdf_utils.cpp:47-63``.
* Self-collision pairs: ``src/fcl_model.cpp:282-290`` and SRDF pruning ``:113-136``.
"""
from __future__ import annotations

import math
import os
import xml.etree.ElementTree as ET
from dataclasses import dataclass, field

import numpy as np

JT_REVOLUTE = 0  # bounded revolute (Pinocchio JointModelR{X,Y,Z} / RevoluteUnaligned)
JT_PRISMATIC = 1  # JointModelP{X,Y,Z} / PrismaticUnaligned
JT_CONTINUOUS = 2  # JointModelRUB{X,Y,Z} / RevoluteUnboundedUnaligned


def quat_to_mat(w, x, y, z):
    """Eigen ``Quaternion::toRotationMatrix`` (no normalisation, like Eigen)."""
    tx, ty, tz = 2 * x, 2 * y, 2 * z
    twx, twy, twz = tx * w, ty * w, tz * w
    txx, txy, txz = tx * x, ty * x, tz * x
    tyy, tyz, tzz = ty * y, tz * y, tz * z
    return np.array(
        [
            [1 - (tyy + tzz), txy - twz, txz + twy],
            [txy + twz, 1 - (txx + tzz), tyz - twx],
            [txz - twy, tyz + twx, 1 - (txx + tyy)],
        ],
        dtype=np.float64,
    )


def mat_to_quat(R):
    """Eigen ``Quaternion(Matrix3)`` -> (w, x, y, z) (Shepperd branches as in Eigen)."""
    t = R[0, 0] + R[1, 1] + R[2, 2]
    if t > 0:
        t = math.sqrt(t + 1.0)
        w = 0.5 * t
        t = 0.5 / t
        return np.array([w, (R[2, 1] - R[1, 2]) * t, (R[0, 2] - R[2, 0]) * t, (R[1, 0] - R[0, 1]) * t])
    i = 0
    if R[1, 1] > R[0, 0]:
        i = 1
    if R[2, 2] > R[i, i]:
        i = 2
    j = (i + 1) % 3
    k = (j + 1) % 3
    t = math.sqrt(R[i, i] - R[j, j] - R[k, k] + 1.0)
    q = np.zeros(4)
    q[1 + i] = 0.5 * t
    t = 0.5 / t
    q[0] = (R[k, j] - R[j, k]) * t
    q[1 + j] = (R[j, i] + R[i, j]) * t
    q[1 + k] = (R[k, i] + R[i, k]) * t
    return q


def rpy_to_quat(roll, pitch, yaw):
    """urdfdom ``Rotation::setFromRPY`` -> (w, x, y, z)."""
    phi, the, psi = roll / 2.0, pitch / 2.0, yaw / 2.0
    x = math.sin(phi) * math.cos(the) * math.cos(psi) - math.cos(phi) * math.sin(the) * math.sin(psi)
    y = math.cos(phi) * math.sin(the) * math.cos(psi) + math.sin(phi) * math.cos(the) * math.sin(psi)
    z = math.cos(phi) * math.cos(the) * math.sin(psi) - math.sin(phi) * math.sin(the) * math.cos(psi)
    w = math.cos(phi) * math.cos(the) * math.cos(psi) + math.sin(phi) * math.sin(the) * math.sin(psi)
    n = math.sqrt(x * x + y * y + z * z + w * w)
    return np.array([w / n, x / n, y / n, z / n])


def pose_to_T(xyz, rpy):
    T = np.eye(4)
    T[:3, :3] = quat_to_mat(*rpy_to_quat(*rpy))
    T[:3, 3] = xyz
    return T


def posevec_to_T(pose7):
    """[x,y,z,qw,qx,qy,qz] -> 4x4 (``src/math_utils.cpp:12-18``)."""
    p = np.asarray(pose7, dtype=np.float64)
    T = np.eye(4)
    T[:3, :3] = quat_to_mat(p[3], p[4], p[5], p[6])
    T[:3, 3] = p[:3]
    return T


def T_to_posevec(T):
    q = mat_to_quat(T[:3, :3])
    return np.array([T[0, 3], T[1, 3], T[2, 3], q[0], q[1], q[2], q[3]])


def _floats(s, n, default):
    if s is None:
        return list(default)
    v = [float(t) for t in s.split()]
    if len(v) != n:
        raise ValueError(f"expected {n} floats, got {s!r}")
    return v


def _origin(elem):
    o = elem.find("origin") if elem is not None else None
    if o is None:
        return np.eye(4)
    return pose_to_T(_floats(o.get("xyz"), 3, (0, 0, 0)), _floats(o.get("rpy"), 3, (0, 0, 0)))


@dataclass
class UrdfCollision:
    gtype: str  # "mesh" | "box" | "sphere" | "cylinder"
    origin: np.ndarray  # 4x4 origin2link
    filename: str = ""
    scale: tuple = (1.0, 1.0, 1.0)
    dims: tuple = ()


@dataclass
class UrdfLink:
    name: str
    collisions: list = field(default_factory=list)
    parent_joint: "UrdfJoint | None" = None
    child_joints: list = field(default_factory=list)


@dataclass
class UrdfJoint:
    name: str
    jtype: str
    parent: str
    child: str
    origin: np.ndarray
    axis: np.ndarray
    lower: float = 0.0
    upper: float = 0.0
    has_limit: bool = False


class UrdfTree:
    """Minimal urdfdom-equivalent tree: links, joints, root, name-sorted children."""

    def __init__(self, xml_string: str):
        try:
            root = ET.fromstring(xml_string)
        except ET.ParseError as e:  # reference: parseURDF returns null -> invalid_argument
            raise ValueError("The XML stream does not contain a valid URDF model.") from e
        if root.tag != "robot":
            raise ValueError("The XML stream does not contain a valid URDF model.")
        self.name = root.get("name", "")
        self.links: dict[str, UrdfLink] = {}
        self.joints: dict[str, UrdfJoint] = {}
        for le in root.findall("link"):
            link = UrdfLink(le.get("name"))
            for ce in le.findall("collision"):
                g = ce.find("geometry")
                if g is None or len(g) == 0:
                    continue
                ge = g[0]
                org = _origin(ce)
                if ge.tag == "mesh":
                    link.collisions.append(
                        UrdfCollision("mesh", org, ge.get("filename"), tuple(_floats(ge.get("scale"), 3, (1, 1, 1))))
                    )
                elif ge.tag == "box":
                    link.collisions.append(UrdfCollision("box", org, dims=tuple(_floats(ge.get("size"), 3, (0, 0, 0)))))
                elif ge.tag == "sphere":
                    link.collisions.append(UrdfCollision("sphere", org, dims=(float(ge.get("radius")),)))
                elif ge.tag == "cylinder":
                    link.collisions.append(
                        UrdfCollision("cylinder", org, dims=(float(ge.get("radius")), float(ge.get("length"))))
                    )
                else:
                    raise ValueError("Unknown geometry type :")
            self.links[link.name] = link
        for je in root.findall("joint"):
            ax = je.find("axis")
            axis = np.array(_floats(ax.get("xyz") if ax is not None else None, 3, (1, 0, 0)))
            lim = je.find("limit")
            j = UrdfJoint(
                je.get("name"),
                je.get("type"),
                je.find("parent").get("link"),
                je.find("child").get("link"),
                _origin(je),
                axis,
            )
            if lim is not None:
                j.has_limit = True
                j.lower = float(lim.get("lower", 0.0))
                j.upper = float(lim.get("upper", 0.0))
            self.joints[j.name] = j
        for jn in sorted(self.joints):  # std::map iteration order
            j = self.joints[jn]
            if j.parent not in self.links or j.child not in self.links:
                raise ValueError(f"joint {jn} references an unknown link")
            self.links[j.child].parent_joint = j
            self.links[j.parent].child_joints.append(j)
        roots = [l for l in self.links.values() if l.parent_joint is None]
        if len(roots) != 1:
            raise ValueError("The XML stream does not contain a valid URDF model.")
        self.root = roots[0]


@dataclass
class KinematicModel:
    """Pinocchio-shaped tables.  Joint index 0 is the universe."""

    joint_names: list  # movable joints, model order (index i <-> model joint i+1)
    joint_parent: np.ndarray  # [nj] model index of parent joint (0 = universe)
    joint_placement: np.ndarray  # [nj,4,4]
    joint_type: np.ndarray  # [nj] JT_*
    joint_axis: np.ndarray  # [nj,3] unit axis
    joint_lower: np.ndarray
    joint_upper: np.ndarray
    joint_short: list  # Pinocchio shortname(), e.g. "JointModelRZ"
    link_names: list  # BODY frames in frame order
    link_joint: np.ndarray  # [nl] model joint index the frame hangs on (0 = universe)
    link_placement: np.ndarray  # [nl,4,4]
    link_parent_link: list  # URDF parent link name ("root's parent" for the root)
    leaf_links: list

    @property
    def nv(self):
        return len(self.joint_names)


def _short_name(jtype, axis):
    ax = None
    for k, u in enumerate(np.eye(3)):
        if np.array_equal(axis, u):
            ax = "XYZ"[k]
    if jtype == JT_REVOLUTE:
        return f"JointModelR{ax}" if ax else "JointModelRevoluteUnaligned"
    if jtype == JT_CONTINUOUS:
        return f"JointModelRUB{ax}" if ax else "JointModelRevoluteUnboundedUnaligned"
    return f"JointModelP{ax}" if ax else "JointModelPrismaticUnaligned"


def build_kinematic_model(tree: UrdfTree) -> KinematicModel:
    jn, jp, jpl, jt, ja, jlo, jhi, jshort = [], [], [], [], [], [], [], []
    ln, lj, lpl, lpar, leaves = [], [], [], [], []
    # frame of a link = (model joint index, placement in that joint's frame)
    frame_of: dict[str, tuple] = {}

    def add_body(name, joint_idx, placement, parent_name):
        ln.append(name)
        lj.append(joint_idx)
        lpl.append(placement)
        lpar.append(parent_name)
        frame_of[name] = (joint_idx, placement)

    add_body(tree.root.name, 0, np.eye(4), "root's parent")

    def dfs(link: UrdfLink):
        j = link.parent_joint
        pj, pplace = frame_of[j.parent]
        if j.jtype == "fixed":
            add_body(link.name, pj, pplace @ j.origin, j.parent)
        elif j.jtype in ("revolute", "continuous", "prismatic"):
            if j.jtype in ("revolute", "prismatic") and not j.has_limit:
                raise RuntimeError(("REVOLUTE" if j.jtype == "revolute" else "PRISMATIC") + " without limits")
            t = {"revolute": JT_REVOLUTE, "continuous": JT_CONTINUOUS, "prismatic": JT_PRISMATIC}[j.jtype]
            n = np.linalg.norm(j.axis)
            if n == 0:
                raise ValueError(f"joint {j.name} has a zero axis")
            jn.append(j.name)
            jp.append(pj)
            jpl.append(pplace @ j.origin)
            jt.append(t)
            ja.append(j.axis / n)
            jshort.append(_short_name(t, j.axis))
            if t == JT_CONTINUOUS:
                jlo.append(-math.inf)
                jhi.append(math.inf)
            else:
                jlo.append(j.lower)
                jhi.append(j.upper)
            add_body(link.name, len(jn), np.eye(4), j.parent)
        else:
            raise ValueError(f"The type of joint {j.name} is not supported.")
        for cj in link.child_joints:
            dfs(tree.links[cj.child])
        if not link.child_joints:
            leaves.append(link.name)

    for cj in tree.root.child_joints:
        dfs(tree.links[cj.child])

    nj = len(jn)
    return KinematicModel(
        jn,
        np.array(jp, dtype=np.int32).reshape(nj),
        np.array(jpl, dtype=np.float64).reshape(nj, 4, 4),
        np.array(jt, dtype=np.int32).reshape(nj),
        np.array(ja, dtype=np.float64).reshape(nj, 3),
        np.array(jlo, dtype=np.float64),
        np.array(jhi, dtype=np.float64),
        jshort,
        ln,
        np.array(lj, dtype=np.int32),
        np.array(lpl, dtype=np.float64).reshape(len(ln), 4, 4),
        lpar,
        leaves,
    )


def collision_dfs(tree: UrdfTree):
    """(link_name, parent_link_name, UrdfCollision) in ``FCLModelTpl::dfs_parse_tree`` order."""
    out = []

    def dfs(link, parent_name):
        for c in link.collisions:
            out.append((link.name, parent_name, c))
        for cj in link.child_joints:
            dfs(tree.links[cj.child], link.name)

    dfs(tree.root, "root's parent")
    return out


def self_collision_pairs(link_names, parent_names, use_parent_rule=True):
    """``src/fcl_model.cpp:282-290`` (with parent/child rule) or ``:61-64`` (without)."""
    pairs = []
    for i in range(len(link_names)):
        for j in range(i):
            if link_names[i] == link_names[j]:
                continue
            if use_parent_rule and (parent_names[i] == link_names[j] or parent_names[j] == link_names[i]):
                continue
            pairs.append((j, i))
    return pairs


def srdf_disabled_pairs(srdf_string: str):
    """``<disable_collisions link1 link2>`` children of ``<robot>`` (``src/fcl_model.cpp:113-136``)."""
    root = ET.fromstring(srdf_string)
    if root.tag != "robot":
        raise RuntimeError("SRDF has no <robot> root")
    return [(n.get("link1"), n.get("link2")) for n in root if n.tag == "disable_collisions"]


def remove_pairs_from_srdf(pairs, link_names, srdf_string):
    dis = srdf_disabled_pairs(srdf_string)
    out = list(pairs)
    for l1, l2 in dis:
        out = [
            (a, b)
            for (a, b) in out
            if not ((link_names[a] == l1 and link_names[b] == l2) or (link_names[a] == l2 and link_names[b] == l1))
        ]
    return out


def read_text(path):
    with open(path, "r") as f:
        return f.read()


def urdf_dir(urdf_filename):
    """``urdf_filename.substr(0, find_last_of("/\\"))`` (``src/fcl_model.cpp:31-32``)."""
    d = os.path.dirname(urdf_filename)
    return d if d else "."
