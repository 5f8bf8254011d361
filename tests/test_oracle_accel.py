"""CPU tests of the oracle's hierarchical leg (the TIMED baseline of bench.py) against the exhaustive leg (the checker):
identical per-pair verdicts, bit for bit, on the C1 / C3 / C5 scenes and on mesh-link robots; plus the world compiler's
unplanned-articulation / attached x attached pair loops."""
import numpy as np

from tests.common import compile_desc, make_panda_world, random_q


def _same(desc, q, threads=8):
    from oracle import oracle as O

    qd = q.astype(np.float32).astype(np.float64)
    v0, p0 = O.collide_batch(desc, qd, threads=threads)
    v1, p1 = O.collide_batch(desc, qd, threads=threads, accel=O.Accel(desc))
    assert np.array_equal(p0, p1), np.argwhere(p0 != p1)[:10]
    assert np.array_equal(v0, v1)
    return p0


def test_accel_equals_exhaustive_c3():
    from tests.workloads import c3

    wl = c3(1024)
    p = _same(wl["desc"], wl["make_q"](0, 10240))
    k = np.array(wl["desc"].pair_kinds)
    assert p[:, k == "articulation_sceneobject"].any() and p[:, k == "attach_sceneobject"].any()
    bt = wl["desc"].obj_type[wl["desc"].pair_b]
    assert p[:, bt == 5].any() and p[:, bt == 6].any(), "mesh and octree pairs must occur in the sample"


def test_accel_equals_exhaustive_c1_c2_c5():
    from tests.workloads import c1_q, c1_world, c2, c5

    for scene in ("A", "B"):
        _same(compile_desc(c1_world(scene)), c1_q(3000))
    wl = c2(1024)
    _same(wl["desc"], wl["make_q"](0, 20000))
    wl = c5(1024)
    _same(wl["desc"], wl["make_q"](0, 6000))


def test_accel_equals_exhaustive_mesh_links():
    """convex=False: BVH links (mesh x mesh self pairs), a mesh obstacle and an octree (mesh x octree)."""
    import os

    from mplib_b200.pymp import fcl
    from tests.common import ROOT
    from tests.workloads import box_surface_points

    w = make_panda_world(convex=False)
    m = fcl.load_mesh_as_BVH(os.path.join(ROOT, "data/panda/franka_description/meshes/collision/link0.stl"))
    w.add_normal_object("mesh_obstacle", fcl.CollisionObject(m, [0.5, 0.2, 0.3], [1, 0, 0, 0]))
    rng = np.random.Generator(np.random.PCG64(4))
    w.add_point_cloud("cloud", box_surface_points(rng, 3000, (0.8, 0.8, 0.05), (0.56, 0, -0.025)), 0.02)
    desc = compile_desc(w)
    p = _same(desc, random_q(600, 17))
    assert p.any()


def test_accel_planner_paths_agree():
    """RRTConnect over the hierarchical checker returns paths the exhaustive motion validator accepts."""
    from mplib_b200.pymp import OMPLPlanner
    from oracle import oracle as O
    from tests.workloads import c4_queries, c4_world

    w = c4_world(cloud_points=5000)
    desc = compile_desc(w)
    ac = O.Accel(desc)
    pl = OMPLPlanner(w)
    lo, hi = pl.lower_joint_limits, pl.upper_joint_limits
    starts, goals = c4_queries(8, 5, lambda q: O.collide_batch(desc, q, threads=8, accel=ac)[0])
    st, paths, gc, sc = O.rrtconnect_batch(desc, lo, hi, starts, goals, range_=0.1, time_limit=5.0, seed=1, threads=8, accel=ac)
    assert (st == 1).sum() >= 4
    for s, p in zip(st, paths):
        if s == 1:
            valid, _ = O.check_motion_batch(desc, lo, hi, p[:-1], p[1:], threads=8)
            assert valid.all()


def test_world_compiler_unplanned_and_attach_attach_pairs():
    from mplib_b200.pymp import ArticulatedModel, PlanningWorld, fcl
    from mplib_b200.workloads import PANDA_HI, PANDA_JOINTS, PANDA_LINKS, PANDA_LO, prefixed_panda_files
    from oracle import oracle as O

    arts = []
    for prefix, xyz, yaw in (("a_", (0.0, -0.3, 0.0), 0.0), ("b_", (0.0, 0.3, 0.0), float(np.pi))):
        urdf, srdf = prefixed_panda_files(prefix, xyz, yaw)
        art = ArticulatedModel(urdf, srdf, (0, 0, -9.81), [prefix + j for j in PANDA_JOINTS], [prefix + l for l in PANDA_LINKS],
                               verbose=False, convex=True)
        art.set_move_group(prefix + "panda_hand")
        arts.append(art)
    w = PlanningWorld(arts, ["robot_a", "robot_b"])
    w.set_articulation_planned("robot_b", False)
    arts[1].set_qpos(np.array([0.3, 0.4, 0.0, -1.6, 0.0, 2.0, 0.8, 0.02, 0.02]), True)
    names = arts[0].get_user_link_names()
    w.attach_object("tool", fcl.Box([0.04, 0.04, 0.25]), "robot_a", names.index("a_panda_hand"), [0, 0, 0.2, 1, 0, 0, 0],
                    ["a_panda_hand", "a_panda_leftfinger", "a_panda_rightfinger"])
    w.attach_object("pack", fcl.Sphere(0.09), "robot_a", names.index("a_panda_link5"), [0.1, 0.0, 0.1, 1, 0, 0, 0],
                    ["a_panda_link5", "a_panda_link4", "a_panda_link6"])
    desc = compile_desc(w)
    k = np.array(desc.pair_kinds)
    # 11 links x 11 unplanned links, 2 attached x 11 unplanned links, 1 attached x attached pair (src/planning_world.cpp:350-367, 395-462)
    assert (k == "articulation_articulation").sum() == 121 and (k == "attach_articulation").sum() == 22 and (k == "attach_attach").sum() == 1
    assert desc.state_dim == 7
    q = random_q(6000, 88, PANDA_LO, PANDA_HI)
    p = _same(desc, q)
    for kind in ("articulation_articulation", "attach_articulation", "attach_attach"):
        assert p[:, k == kind].any(), kind
    # the unplanned robot stays where its current qpos puts it: moving it changes the verdicts
    arts[1].set_qpos(np.array([-0.3, -0.4, 0.0, -1.6, 0.0, 2.0, 0.8, 0.0, 0.0]), True)
    p2 = O.collide_batch(compile_desc(w), q.astype(np.float32).astype(np.float64), threads=8)[1]
    assert (p2 != p).any()
