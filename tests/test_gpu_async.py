"""GPU tests of the round-2 entry points: sparse hit list, device-pointer calls that return after enqueueing
(+ mpb_scene_poll), in-place object updates, the small-batch CUDA-graph path, and the two advisor scenarios
(shared-memory limit across scenes; fixed x fixed pairs in distance queries)."""
import ctypes as C

import numpy as np
import pytest

from tests.common import make_panda_world, random_q

pytestmark = pytest.mark.gpu


def _c3_small_world(n_each=8, mesh=True, cloud_points=5000):
    from tests.workloads import c3_world

    return c3_world(n_each=n_each, mesh=mesh, cloud_points=cloud_points)


def test_hit_list_equals_dense_mask():
    w = _c3_small_world()
    dev = w._sync()["dev"]
    for n in (1, 37, 3000, 70000):  # graph path, graph path, graph path, chunked path
        q = random_q(n, 300 + n)
        v, ph = dev.collide(q, want_pairs=True)
        vh, hits = dev.collide_hits(q)
        assert np.array_equal(v, vh)
        dense = np.zeros_like(ph)
        dense[hits[:, 0], hits[:, 1]] = True
        assert np.array_equal(dense, ph)
        assert len(np.unique(hits, axis=0)) == len(hits)  # one entry per (configuration, pair)
        v2, ph2 = dev.collide(q, want_pairs=True)
        assert np.array_equal(ph, ph2), "two calls on the same input must agree bit for bit"
    # a capacity that is too small truncates the list and reports the full count
    q = random_q(3000, 9)
    _, full = dev.collide_hits(q)
    assert len(full) > 16
    vb = np.zeros((3000 + 31) // 32, dtype=np.uint32)
    hits = np.zeros((16, 2), dtype=np.uint32)
    nh = np.zeros(1, dtype=np.uint32)
    dev.collide_hits_raw(np.ascontiguousarray(q.astype(np.float32)), 3000, vb, hits, 16, nh)
    assert int(nh[0]) == len(full)
    assert set(map(tuple, hits.tolist())) <= set(map(tuple, full.tolist()))


def test_device_pointer_calls_are_async_and_polled():
    import torch

    from mplib_b200 import _lib

    w = _c3_small_world()
    dev = w._sync()["dev"]
    n = 50000
    qs = [random_q(n, 500 + k).astype(np.float32) for k in range(4)]
    qd = [torch.from_numpy(q).cuda() for q in qs]
    W = dev.pair_words
    vd = [torch.zeros((n + 31) // 32, dtype=torch.int32, device="cuda") for _ in qs]
    pd = [torch.zeros((n, W), dtype=torch.int32, device="cuda") for _ in qs]
    stream = torch.cuda.Stream()
    sts = [dev.collide_raw(qd[k], n, vd[k], pd[k], _lib.MPB_DEVICE, stream.cuda_stream) for k in range(4)]
    # the first device-pointer call calibrates through the blocking protocol, the others return after enqueueing
    assert sts[0]["async_pending"] == 0 and all(s["async_pending"] == 1 for s in sts[1:])
    acc, done, failed = dev.poll(wait=True)
    assert done == 3 and failed == 0 and acc["n_states"] == 3 * n and acc["narrow_iterations"] > 0
    torch.cuda.synchronize()
    ref = [dev.collide(q, want_pairs=True) for q in qs]
    # a larger batch than any calibrated one goes through the blocking protocol again, a smaller one does not
    big = torch.from_numpy(random_q(2 * n, 1).astype(np.float32)).cuda()
    vb = torch.zeros((2 * n + 31) // 32, dtype=torch.int32, device="cuda")
    assert dev.collide_raw(big, 2 * n, vb, None, _lib.MPB_DEVICE, stream.cuda_stream)["async_pending"] == 0
    assert dev.collide_raw(big, n // 2, vb, None, _lib.MPB_DEVICE, stream.cuda_stream)["async_pending"] == 1
    dev.poll(wait=True)
    for k in range(4):
        v = np.unpackbits(vd[k].cpu().numpy().view(np.uint8), bitorder="little")[:n].astype(bool)
        p = np.unpackbits(pd[k].cpu().numpy().view(np.uint8), axis=1, bitorder="little")[:, : dev.n_pairs].astype(bool)
        assert np.array_equal(v, ref[k][0]) and np.array_equal(p, ref[k][1])
    # hit-list flavour, device pointers, with per-pass timing switched on
    dev.set_async_timing(True)
    hits = torch.zeros((8 * n, 2), dtype=torch.int32, device="cuda")
    nh = torch.zeros(1, dtype=torch.int32, device="cuda")
    st = dev.collide_hits_raw(qd[1], n, vd[1], hits, 8 * n, nh, _lib.MPB_DEVICE, stream.cuda_stream)
    assert st["async_pending"] == 1
    acc, done, failed = dev.poll(wait=True)
    assert done == 1 and failed == 0 and acc["gpu_ms"] > 0
    k = int(nh.item())
    h = hits[:k].cpu().numpy().astype(np.int64)
    dense = np.zeros_like(ref[1][1])
    dense[h[:, 0], h[:, 1]] = True
    assert np.array_equal(dense, ref[1][1])
    # a blocking host call while nothing is pending, then an empty poll
    dev.collide(qs[0][:10])
    acc, done, failed = dev.poll(wait=False)
    assert done == 0 and failed == 0


def test_update_objects_in_place_matches_recompile():
    from mplib_b200.pymp import fcl
    from oracle import oracle as O

    def build(shift):
        w = make_panda_world()
        w.add_normal_object("box", fcl.CollisionObject(fcl.Box([0.2, 0.3, 0.25]), [0.45 + shift, 0.1, 0.35], [1, 0, 0, 0]))
        w.add_normal_object("ball", fcl.CollisionObject(fcl.Sphere(0.12), [0.3, -0.3 - shift, 0.5], [1, 0, 0, 0]))
        return w

    q = random_q(6000, 41)
    w = build(0.0)
    c0 = w._sync()
    v0, p0 = c0["dev"].collide(q)
    # move both obstacles the way SapienPlanningWorld.update_from_simulation does (set_transformation in place)
    w.normal_objects["box"].set_transformation([0.55, 0.1, 0.35, 0.9238795, 0, 0, 0.3826834])
    w.normal_objects["ball"].set_transformation([0.3, -0.4, 0.5, 1, 0, 0, 0])
    c1 = w._sync()
    assert c1["dev"] is c0["dev"], "a pose change must not re-compile the scene"
    v1, p1 = c1["dev"].collide(q)
    ref = make_panda_world()
    ref.add_normal_object("box", fcl.CollisionObject(fcl.Box([0.2, 0.3, 0.25]), [0.55, 0.1, 0.35], [0.9238795, 0, 0, 0.3826834]))
    ref.add_normal_object("ball", fcl.CollisionObject(fcl.Sphere(0.12), [0.3, -0.4, 0.5], [1, 0, 0, 0]))
    v2, p2 = ref._sync()["dev"].collide(q)
    assert np.array_equal(v1, v2) and np.array_equal(p1, p2)
    assert (v0 != v1).any()
    # the description the oracle sees follows the update
    ov, op = O.collide_batch(c1["dev"].desc, q.astype(np.float32).astype(np.float64), threads=8)
    assert (op != p1).sum() <= 4
    # attached body re-posed in place
    from tests.workloads import attach_c3_box

    attach_c3_box(w)
    ca = w._sync()
    va, _ = ca["dev"].collide(q)
    w.attached_bodies["grasped"].set_pose(np.array([[1, 0, 0, 0.0], [0, 1, 0, 0.0], [0, 0, 1, 0.30], [0, 0, 0, 1.0]]))
    cb = w._sync()
    assert cb["dev"] is ca["dev"]
    vb, pb = cb["dev"].collide(q)
    ob, opb = O.collide_batch(cb["dev"].desc, q.astype(np.float32).astype(np.float64), threads=8)
    assert (opb != pb).sum() <= 4 and (va != vb).any()


def test_small_batch_graph_path_is_consistent():
    """Calls of <= 4096 states replay a cached CUDA graph: same results as the chunked path, across scene edits and
    q_base changes (legacy single-state calls change the articulation's qpos before every query)."""
    w = _c3_small_world(n_each=4, mesh=False, cloud_points=2000)
    dev = w._sync()["dev"]
    q = random_q(5000, 77)
    v_big, p_big = dev.collide(q)  # chunked path
    for n in (1, 2, 33, 4096):
        for rep in range(3):  # the second and third call replay the graph
            v, p = dev.collide(q[:n])
            assert np.array_equal(v, v_big[:n]) and np.array_equal(p, p_big[:n])
    # legacy stateful calls: a finger-joint change (outside the move group) must reach the device
    art = w.get_articulation("robot")
    full = np.array([0, 0.19, 0.0, -2.61, 0.0, 2.94, 0.78, 0.0, 0.0])
    art.set_qpos(full, True)
    a = w.collide()
    full[7:] = 0.04
    art.set_qpos(full, True)
    b = w.collide()
    assert isinstance(a, bool) and isinstance(b, bool)
    got = w._sync()["dev"].collide(full[None, :7])[0][0]
    assert got == b
    # ACM edit invalidates the captured pair tables
    acm = w.get_allowed_collision_matrix()
    names = [n for n in w.normal_objects if n.startswith("sphere")]
    v1, _ = w._sync()["dev"].collide(q[:100])
    for nme in names:
        acm.set_default_entry(nme, True)
    v2, p2 = w._sync()["dev"].collide(q[:100])
    v3, p3 = w._sync()["dev"].collide(np.concatenate([q[:100], q[:5000]]))  # chunked path on the same pair list
    assert np.array_equal(v2, v3[:100]) and np.array_equal(p2, p3[:100])


def test_second_scene_does_not_lower_shared_memory_limit():
    """world collide, then a geometry-less PinocchioModel scene (FK / IK), then collide again."""
    from mplib_b200.workloads import make_panda

    w = make_panda_world()
    q = random_q(3000, 3)
    v0, _ = w._sync()["dev"].collide(q)
    art = make_panda()
    pm = art.get_pinocchio_model()
    pm.compute_forward_kinematics(np.zeros(9))
    pose = pm.get_link_pose(9)
    pm.compute_IK_CLIK(9, pose, np.zeros(9))
    v1, _ = w._sync()["dev"].collide(q)
    assert np.array_equal(v0, v1)


def test_distance_with_fixed_fixed_pair_nearest():
    """the base link sits on the world joint, so panda_link0 x obstacle is a fixed x fixed pair; when it is
    the nearest pair the distance query must report it."""
    from mplib_b200.pymp import fcl
    from oracle import oracle as O

    w = make_panda_world()
    w.add_normal_object("near_base", fcl.CollisionObject(fcl.Sphere(0.05), [-0.30, 0.0, 0.05], [1, 0, 0, 0]))
    q = random_q(500, 12)
    c = w._sync()
    md, am = c["dev"].distance(q)
    omd, oam = O.distance_batch(c["dev"].desc, q.astype(np.float32).astype(np.float64))
    kinds = c["dev"].desc
    base_pair = [k for k, p in enumerate(c["pairs"]) if p[6] == "panda_link0" and p[7] == "near_base"]
    assert len(base_pair) == 1
    assert (oam == base_pair[0]).sum() > 10, "the scenario must make the fixed x fixed pair the nearest one for some states"
    ok = (md < 0) == (omd < 0)
    assert ok.mean() > 0.99
    sel = (md > 0) & (omd > 0)
    assert np.abs(md[sel] - omd[sel]).max() < 1e-4
    assert (am[sel & (oam == base_pair[0])] == base_pair[0]).mean() > 0.95
    # a world whose ONLY admitted pair is fixed x fixed
    c["dev"].set_pairs([c["pairs"][base_pair[0]][0]], [c["pairs"][base_pair[0]][1]])
    md1, am1 = c["dev"].distance(q[:8])
    o1, _ = O.distance_batch(c["dev"].desc, q[:8].astype(np.float32).astype(np.float64))
    assert np.abs(md1 - o1).max() < 1e-4 and (am1 == 0).all()


def test_unplanned_articulation_and_two_attached_bodies():
    """collideWithOthers over an UNPLANNED articulation (articulation_articulation, attach_articulation) and two attached
    bodies (attach_attach): src/planning_world.cpp:350-367, 395-417, 437-462."""
    from mplib_b200.pymp import ArticulatedModel, PlanningWorld, fcl
    from mplib_b200.workloads import PANDA_JOINTS, PANDA_LINKS, prefixed_panda_files
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
    hand = arts[0].get_user_link_names().index("a_panda_hand")
    l5 = arts[0].get_user_link_names().index("a_panda_link5")
    w.attach_object("tool", fcl.Box([0.04, 0.04, 0.25]), "robot_a", hand, [0, 0, 0.2, 1, 0, 0, 0],
                    ["a_panda_hand", "a_panda_leftfinger", "a_panda_rightfinger"])
    w.attach_object("pack", fcl.Sphere(0.09), "robot_a", l5, [0.1, 0.0, 0.1, 1, 0, 0, 0], ["a_panda_link5", "a_panda_link4", "a_panda_link6"])
    c = w._sync()
    desc = c["dev"].desc
    kinds = set(desc.pair_kinds)
    assert {"articulation_articulation", "attach_articulation", "attach_attach", "self", "self_attach"} <= kinds, kinds
    assert desc.state_dim == 7  # only robot_a is planned
    from mplib_b200.workloads import PANDA_HI, PANDA_LO

    q = random_q(6000, 88, PANDA_LO, PANDA_HI)
    v, ph = c["dev"].collide(q)
    qd = q.astype(np.float32).astype(np.float64)
    ov, oph = O.collide_batch(desc, qd, threads=8)
    mism = np.argwhere(ph != oph)
    for i, p in mism:
        assert abs(O.pair_signed_distance(desc, qd[i], int(p))) < 1e-4
    assert len(mism) <= 6
    k = np.array(desc.pair_kinds)
    for kind in ("articulation_articulation", "attach_articulation", "attach_attach"):
        assert oph[:, k == kind].any(), f"no {kind} collision in the sample: the test would not exercise that loop"
    # legacy query reports the types
    hit = np.argwhere(oph[:, k == "attach_attach"].any(axis=1)).ravel()
    w.set_qpos_all(q[hit[0]])
    assert any(r.collision_type == "attach_attach" for r in w.collide_full())


def test_native_module_and_ctypes_binding_agree(monkeypatch):
    """The per-call path through the pybind11 module (default) and through the ctypes fallback binding: same results."""
    from mplib_b200 import _lib

    assert _lib._native is not None
    w = _c3_small_world(n_each=4)
    desc = w._sync()["dev"].desc
    q = random_q(3000, 5)
    nat = _lib.DeviceScene(desc, 0)
    assert nat._n is not None
    monkeypatch.setattr(_lib, "_native", None)
    ct = _lib.DeviceScene(desc, 0)
    assert ct._n is None
    for n in (1, 3000):
        a, b = nat.collide(q[:n], stats=True), ct.collide(q[:n], stats=True)
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and a[2]["n_states"] == b[2]["n_states"] == n
        a, b = nat.collide_hits(q[:n]), ct.collide_hits(q[:n])
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
        assert np.array_equal(nat.fk(q[:n]), ct.fk(q[:n]))
        a, b = nat.distance(q[:n]), ct.distance(q[:n])
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    v, _ = nat.collide(q[:10], want_pairs=False)
    assert _ is None and v.dtype == bool
    with pytest.raises(RuntimeError, match="State dimension"):
        nat.collide(np.zeros((3, 5)))
    nat.close(); ct.close()


def test_multi_pass_host_call_matches_single_pass_calls(panda_world):
    """A host-buffer call larger than one pass (2^20 configurations) double-buffers its passes and copies every pass's share
    of the hit list out on a stream of its own under the next pass's kernels: same verdicts, pair masks and hit list as the
    same configurations sent in single-pass calls."""
    dev = panda_world._sync()["dev"]
    n = (1 << 20) + 70001
    q = random_q(n, 77)
    v, ph = dev.collide(q, want_pairs=True)
    for lo, hi in ((0, 300000), (700000, 1000000), (n - 200000, n)):
        v2, ph2 = dev.collide(q[lo:hi], want_pairs=True)
        assert np.array_equal(v[lo:hi], v2) and np.array_equal(ph[lo:hi], ph2)
    vh, hits = dev.collide_hits(q)
    assert np.array_equal(vh, v)
    got = np.zeros_like(ph)
    got[hits[:, 0], hits[:, 1]] = True
    assert np.array_equal(got, ph) and len(hits) == int(ph.sum())
