"""scene ingestion formats.  `ArticulatedModel.create_from_urdf_string` (URDF without geometry +
explicit per-link collision objects, src/articulated_model.cpp:38-67, src/fcl_model.cpp:37-67) and the data-level form
of SapienPlanningWorld's shape conversion (mplib/sapien_utils/conversion.py:326-409)."""
import os
import xml.etree.ElementTree as ET

import numpy as np
import pytest

from tests.common import PANDA_JOINTS, PANDA_LINKS, SRDF, URDF, compile_desc, make_panda_world, random_q


def _panda_from_string():
    """The Panda of tests/common.py rebuilt the SAPIEN way: geometry stripped from the URDF, one convex_mesh record
    per <collision> with the hull's vertices and triangles at the collision origin."""
    from mplib_b200 import urdf as U
    from mplib_b200.pymp import PlanningWorld, fcl
    from mplib_b200.sapien_utils import articulation_from_link_shapes

    text = U.read_text(URDF)
    tree = U.UrdfTree(text)
    shapes = {}
    for link_name, _, c in U.collision_dfs(tree):
        hull = fcl.load_mesh_as_Convex(os.path.join(U.urdf_dir(URDF), c.filename + ".convex.stl"), c.scale)
        rec = {"type": "convex_mesh", "vertices": hull.vertices, "triangles": hull.faces, "scale": (1.0, 1.0, 1.0),
               "local_pose": (c.origin[:3, 3], U.mat_to_quat(c.origin[:3, :3]))}
        shapes.setdefault(link_name, []).append(rec)
    root = ET.fromstring(text)
    for link in root.iter("link"):
        for tag in ("collision", "visual"):
            for el in link.findall(tag):
                link.remove(el)
    bare = ET.tostring(root, encoding="unicode")
    art = articulation_from_link_shapes(bare, U.read_text(SRDF), [(n, shapes.get(n, [])) for n in PANDA_LINKS],
                                        joint_names=PANDA_JOINTS, link_names=PANDA_LINKS, name="robot")
    art.set_move_group("panda_hand")
    return art, PlanningWorld([art], ["robot"])


def test_shape_records_convert_like_the_reference():
    from mplib_b200.pymp import fcl
    from mplib_b200.sapien_utils import convert_col_shapes

    objs = convert_col_shapes([{"type": "capsule", "radius": 0.05, "half_length": 0.2, "local_pose": ((0.1, 0, 0), (1, 0, 0, 0))}],
                              entity_pose=((0, 0, 1.0), (np.cos(0.3), 0, 0, np.sin(0.3))), name="pole")
    assert len(objs) == 1 and isinstance(objs[0].get_collision_geometry(), fcl.Capsule)
    g = objs[0].get_collision_geometry()
    assert g.radius == 0.05 and g.lz == 0.4
    R = objs[0].get_rotation()
    # the capsule's FCL z axis is the PhysX x axis of the shape, rotated by the entity's yaw of 0.6 rad
    assert np.allclose(R[:, 2], [np.cos(0.6), np.sin(0.6), 0.0]) and np.allclose(objs[0].get_translation(), [0.1 * np.cos(0.6), 0.1 * np.sin(0.6), 1.0])
    # an articulation link ignores the entity pose; planes are skipped; triangle meshes are refused
    link_objs = convert_col_shapes([{"type": "box", "half_size": (0.1, 0.2, 0.3)}, {"type": "plane"}], entity_pose=((5, 5, 5), (1, 0, 0, 0)),
                                   is_articulation_link=True)
    assert len(link_objs) == 1 and np.allclose(link_objs[0].get_translation(), 0) and np.allclose(link_objs[0].get_collision_geometry().side, [0.2, 0.4, 0.6])
    with pytest.raises(NotImplementedError):
        convert_col_shapes([{"type": "triangle_mesh"}])
    with pytest.raises(TypeError):
        convert_col_shapes([{"type": "torus"}])
    assert convert_col_shapes([]) == []


def test_urdf_string_panda_equals_the_file_panda_on_the_oracle():
    from oracle import oracle as O

    art, world = _panda_from_string()
    ref_world = make_panda_world()
    fm, rm = art.get_fcl_model(), ref_world.get_articulation("robot").get_fcl_model()
    assert fm.get_collision_link_names() == rm.get_collision_link_names()
    # no parent/child rule on this path (src/fcl_model.cpp:61-64): the SRDF alone leaves the same 19 pairs
    assert fm.get_collision_pairs() == rm.get_collision_pairs() and len(fm.get_collision_pairs()) == 19
    q = random_q(2000, 5)
    a = O.collide_batch(compile_desc(world), q)
    b = O.collide_batch(compile_desc(ref_world), q)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and 0.01 < a[0].mean() < 0.95


@pytest.mark.gpu
def test_urdf_string_panda_on_the_device():
    """The string-built world gives the same verdict words as the file-built one (same geometry, same pair list), and
    a link set made of primitive shape records agrees with the oracle outside the 1e-4 m contact band."""
    from oracle import oracle as O
    from mplib_b200 import urdf as U
    from mplib_b200.pymp import PlanningWorld
    from mplib_b200.sapien_utils import articulation_from_link_shapes
    from tests.test_gpu_parity import _check_pairs

    art, world = _panda_from_string()
    ref_world = make_panda_world()
    q = random_q(20000, 6)
    v, p = world.collide_full_batch(q)
    rv, rp = ref_world.collide_full_batch(q)
    assert np.array_equal(v, rv) and np.array_equal(p, rp) and 0.01 < v.mean() < 0.95

    # capsules along the links (PhysX x axis -> FCL z axis), a cylinder forearm, a box hand
    cap = lambda r, h, p=(0, 0, 0): {"type": "capsule", "radius": r, "half_length": h, "local_pose": (p, (np.cos(np.pi / 4), 0, -np.sin(np.pi / 4), 0))}
    shapes = [("panda_link0", [{"type": "box", "half_size": (0.1, 0.1, 0.07), "local_pose": ((0, 0, 0.07), (1, 0, 0, 0))}]),
              ("panda_link2", [cap(0.06, 0.08)]), ("panda_link4", [cap(0.06, 0.08)]),
              ("panda_link5", [{"type": "cylinder", "radius": 0.05, "half_length": 0.12, "local_pose": ((0, 0, -0.15), (np.cos(np.pi / 4), 0, -np.sin(np.pi / 4), 0))}]),
              ("panda_link7", [{"type": "sphere", "radius": 0.06}]),
              ("panda_hand", [{"type": "box", "half_size": (0.03, 0.1, 0.04), "local_pose": ((0, 0, 0.03), (1, 0, 0, 0))}])]
    text = U.read_text(URDF)
    near = [("panda_link0", "panda_link2"), ("panda_link2", "panda_link4"), ("panda_link4", "panda_link5"), ("panda_link5", "panda_link7"),
            ("panda_link7", "panda_hand")]
    srdf = "<robot name='panda'>" + "".join(f"<disable_collisions link1='{a}' link2='{b}' reason='Adjacent'/>" for a, b in near) + "</robot>"
    art2 = articulation_from_link_shapes(text, srdf, shapes, joint_names=PANDA_JOINTS, link_names=PANDA_LINKS, name="robot")
    art2.set_move_group("panda_hand")
    w2 = PlanningWorld([art2], ["robot"])
    desc = compile_desc(w2)
    assert desc.n_pairs == 10  # 6 objects: 15 pairs (same-link exclusion only) minus the 5 SRDF entries
    v2, p2 = w2.collide_full_batch(q)
    qd = q.astype(np.float32).astype(np.float64)
    ov, op = O.collide_batch(desc, qd)
    n_mis, _ = _check_pairs(desc, qd, p2, op)
    print(f"primitive links: collide rate {ov.mean():.3f}, mismatches {n_mis} (all in band)")
    assert n_mis <= 20 and 0.02 < ov.mean() < 0.98
