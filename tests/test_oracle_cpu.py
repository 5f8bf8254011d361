"""CPU tests that pin the oracle (oracle/oracle.c) -- the reference cannot run here and its tests hold no
golden vectors (parity unpinned), so the oracle is pinned by
  (1) hand-derived FK known answers from data/panda/panda.urdf,
  (2) the qualitative outcomes stated in examples/detect_collision.py:22-65,
  (3) analytic primitive cases with closed-form signed distances,
  (4) an independent exhaustive SAT on the Panda hulls as a second opinion on the MPR restatement,
  (5) the 55 -> 46 -> 19 pair-list known answer,
  (6) the octomap touched-voxel rule.
"""
import numpy as np
import pytest

from tests.common import PANDA_HI, PANDA_LINKS, PANDA_LO, compile_desc, make_panda_world, random_q


@pytest.fixture(scope="module")
def panda_desc(panda_world):
    return compile_desc(panda_world)


def _quat_R(q):
    from mplib_b200.urdf import quat_to_mat

    return quat_to_mat(*q)


# ------------------------------------------------------------------------------------------- (1) FK
def test_fk_known_answers_zero(panda_desc):
    """q = 0."""
    from oracle import oracle as O

    fk = O.fk_batch(panda_desc, np.zeros((1, 7)))[0]
    p = {n: fk[i] for i, n in enumerate(panda_desc.link_names)}
    assert np.allclose(p["panda_link8"][:3], [0.088, 0, 0.926], atol=1e-9)
    assert np.allclose(_quat_R(p["panda_link8"][3:]), np.diag([1.0, -1.0, -1.0]), atol=1e-9)
    assert np.allclose(p["panda_link7"][:3], [0.088, 0, 1.033], atol=1e-9)
    assert np.allclose(p["panda_hand"][:3], [0.088, 0, 0.926], atol=1e-9)
    assert np.allclose(_quat_R(p["panda_hand"][3:])[0], [0.707107, 0.707107, 0], atol=1e-6)
    for f in ("panda_leftfinger", "panda_rightfinger"):
        assert np.allclose(p[f][:3], [0.088, 0, 0.8676], atol=1e-9)


def test_fk_known_answers_test_basic_pose(panda_world):
    """q = (0,0.2,0,-2.6,0,3.0,0.8,0,0) from tests/test_basic.py:42 of the reference."""
    from oracle import oracle as O

    desc = compile_desc(panda_world)
    fk = O.fk_batch(desc, np.array([[0, 0.2, 0, -2.6, 0, 3.0, 0.8]]))[0]
    p = {n: fk[i] for i, n in enumerate(desc.link_names)}
    assert np.allclose(p["panda_link8"][:3], [0.457507, 0, 0.204750], atol=2e-6)
    R = _quat_R(p["panda_hand"][3:])
    assert np.allclose(R[0], [0.979962, -0.014310, 0.198669], atol=2e-6)
    assert np.allclose(R[2], [0.198648, -0.002901, -0.980067], atol=2e-6)
    for f in ("panda_leftfinger", "panda_rightfinger"):
        assert np.allclose(p[f][:3], [0.469110, 0, 0.147514], atol=2e-6)


def test_fk_independent_dh_chain(panda_desc):
    """Second opinion: the Panda arm as an explicit product of the URDF joint origins (rpy, xyz) and
    Rz(q), written out here independently of the scene compiler (data/panda/panda.urdf:31-157)."""
    from oracle import oracle as O

    def Rx(a):
        c, s = np.cos(a), np.sin(a)
        return np.array([[1, 0, 0, 0], [0, c, -s, 0], [0, s, c, 0], [0, 0, 0, 1.0]])

    def Rz(a):
        c, s = np.cos(a), np.sin(a)
        return np.array([[c, -s, 0, 0], [s, c, 0, 0], [0, 0, 1, 0], [0, 0, 0, 1.0]])

    def Tr(x, y, z):
        T = np.eye(4)
        T[:3, 3] = [x, y, z]
        return T

    h = np.pi / 2
    origins = [Tr(0, 0, 0.333), Rx(-h), Tr(0, -0.316, 0) @ Rx(h), Tr(0.0825, 0, 0) @ Rx(h),
               Tr(-0.0825, 0.384, 0) @ Rx(-h), Rx(h), Tr(0.088, 0, 0) @ Rx(h)]
    q = random_q(64, 77)
    fk = O.fk_batch(panda_desc, q)
    i7 = panda_desc.link_names.index("panda_link7")
    i8 = panda_desc.link_names.index("panda_link8")
    for k in range(q.shape[0]):
        T = np.eye(4)
        for j in range(7):
            T = T @ origins[j] @ Rz(q[k, j])
        assert np.allclose(fk[k, i7, :3], T[:3, 3], atol=1e-12)
        assert np.allclose(_quat_R(fk[k, i7, 3:]), T[:3, :3], atol=1e-12)
        T8 = T @ Tr(0, 0, 0.107)
        assert np.allclose(fk[k, i8, :3], T8[:3, 3], atol=1e-12)


# ------------------------------------------------------------------------------------------- (5) pair list
EXPECTED_19 = {
    ("panda_link0", x) for x in ("panda_link5", "panda_link6", "panda_link7", "panda_hand", "panda_leftfinger", "panda_rightfinger")
} | {
    ("panda_link1", x) for x in ("panda_link5", "panda_link6", "panda_link7", "panda_hand", "panda_leftfinger", "panda_rightfinger")
} | {
    ("panda_link2", x) for x in ("panda_link5", "panda_link7", "panda_hand", "panda_leftfinger", "panda_rightfinger")
} | {("panda_link5", "panda_leftfinger"), ("panda_link5", "panda_rightfinger")}


def test_pair_list_55_46_19(panda_world):
    """src/fcl_model.cpp:282-290 (parent/child rule on the *link* parent) then :113-136 (SRDF)."""
    from mplib_b200 import urdf as U

    fm = panda_world.get_articulation("robot").get_fcl_model()
    names = fm.collision_link_names
    assert len(names) == 11 and "panda_link8" not in names
    all_pairs = U.self_collision_pairs(names, None, False)
    assert len(all_pairs) == 55
    after_parent = U.self_collision_pairs(names, fm.parent_link_names, True)
    assert len(after_parent) == 46
    # hand's link parent is the geometry-less panda_link8, so hand x link7 survives the parent rule
    assert (names.index("panda_link7"), names.index("panda_hand")) in after_parent
    got = [(names[a], names[b]) for a, b in fm.collision_pairs]
    assert len(got) == 19 and set(got) == EXPECTED_19
    # order: (j, i) for i ascending, j < i
    idx = [(a, b) for a, b in fm.collision_pairs]
    assert idx == sorted(idx, key=lambda ab: (ab[1], ab[0]))
    assert all(a < b for a, b in idx)


def test_hull_sizes_match_survey(panda_desc):
    """unique hull vertices per link after aiProcess_JoinIdenticalVertices-style de-dup."""
    nv = np.diff(panda_desc.hull_vert_off)
    by_name = {}
    for o in range(panda_desc.n_objects):
        by_name[panda_desc.object_names[o]] = int(nv[panda_desc.obj_geom[o]])
    expect = {"panda_link0": 84, "panda_link1": 133, "panda_link2": 131, "panda_link3": 131, "panda_link4": 133,
              "panda_link5": 117, "panda_link6": 102, "panda_link7": 79, "panda_hand": 93, "panda_leftfinger": 15,
              "panda_rightfinger": 15}
    assert by_name == expect
    assert sum(expect.values()) == 1033


# ------------------------------------------------------------------------------------------- (2) example outcomes
def test_detect_collision_example_outcomes(panda_world):
    """examples/detect_collision.py:22-59 through the oracle."""
    from oracle import oracle as O
    from mplib_b200.pymp import PlanningWorld
    from tests.common import make_panda
    from tests.workloads import box_surface_points

    desc = compile_desc(panda_world)
    v, ph = O.collide_batch(desc, np.array([[0, 0.19, 0.0, -2.61, 0.0, 2.94, 0.78]]))
    assert not v[0] and not ph.any()
    v, ph = O.collide_batch(desc, np.array([[0, 1.36, 0, -3, -3, 3, -1]]))
    assert v[0] and ph.any()
    # env: 10k-point floor cloud, res 1e-3 (mplib/planner.py:377-378)
    w = PlanningWorld([make_panda()], ["robot"])
    rng = np.random.Generator(np.random.PCG64(100))
    w.add_point_cloud("scene_pcd", box_surface_points(rng, 10000, (2, 2, 0.1), (0, 0, -0.1)), 1e-3)
    d2 = compile_desc(w)
    assert d2.n_pairs == 19 + 11
    v, ph = O.collide_batch(d2, np.array([[0, 1.5, 0, -1.5, 0, 0, 0]]))
    env_cols = [k for k in range(d2.n_pairs) if d2.pair_kinds[k] == "articulation_sceneobject"]
    assert v[0] and ph[0, env_cols].any()
    # the base does not touch the floor (the cloud was shifted down "so we do not collide with the base")
    v, ph = O.collide_batch(d2, np.array([[0, 0.19, 0.0, -2.61, 0.0, 2.94, 0.78]]))
    assert not ph[0, env_cols].any()


def test_ik_clik_round_trip(panda_world):
    """Oracle restatement of computeIKCLIK (src/pinocchio_model.cpp:377-432): pose -> q -> pose round trip through the
    oracle's own FK, masked joints untouched, an out-of-reach pose reported as a failure."""
    from oracle import oracle as O

    desc = compile_desc(panda_world)
    li = desc.link_names.index("panda_hand")
    rng = np.random.Generator(np.random.PCG64(9))
    q_true = PANDA_LO + (PANDA_HI - PANDA_LO) * rng.random((256, 7))
    poses = O.fk_batch(desc, q_true)[:, li, :]
    q0 = q_true + rng.normal(0, 0.2, q_true.shape)
    q, ok, err = O.ik_clik_batch(desc, li, poses, q0)
    assert ok.all() and np.linalg.norm(err, axis=1).max() < 1e-5
    back = O.fk_batch(desc, q)[:, li, :]
    assert np.abs(back[:, :3] - poses[:, :3]).max() < 2e-5
    mask = np.array([1, 0, 0, 0, 0, 0, 0], dtype=np.uint8)
    qm, okm, _ = O.ik_clik_batch(desc, li, poses[:16], q0[:16], mask, max_iter=200)
    assert np.array_equal(qm[:, 0], q0[:16, 0])
    far = poses[:4].copy()
    far[:, :3] = [3.0, 0.0, 0.5]
    _, okf, _ = O.ik_clik_batch(desc, li, far, q0[:4], max_iter=200)
    assert not okf.any()


def test_link_jacobian_matches_finite_differences(panda_world):
    """Oracle restatement of computeFullJacobian + getLinkJacobian (src/pinocchio_model.cpp:335-360) pinned by central
    differences of the oracle's own FK: column j of the WORLD Jacobian is the spatial twist (v at the world origin, w)
    of d/dq_j, i.e. dT/dq_j T^-1 = [[w]x, v]; the LOCAL form is the body twist T^-1 dT/dq_j."""
    from oracle import oracle as O

    desc = compile_desc(panda_world)
    li = desc.link_names.index("panda_hand")
    rng = np.random.Generator(np.random.PCG64(17))
    q = PANDA_LO + (PANDA_HI - PANDA_LO) * rng.random((32, 7))
    Jw, T = O.link_jacobian_batch(desc, li, q, local=False)
    Jl, _ = O.link_jacobian_batch(desc, li, q, local=True)
    assert np.abs(T[:, :, 3] - O.fk_batch(desc, q)[:, li, :3]).max() < 1e-12
    h = 1e-6

    def hom(q_):
        _, T_ = O.link_jacobian_batch(desc, li, q_)
        M = np.tile(np.eye(4), (T_.shape[0], 1, 1))
        M[:, :3, :] = T_
        return M

    T0 = hom(q)
    for j in range(7):
        e = np.zeros(7)
        e[j] = h
        dT = (hom(q + e) - hom(q - e)) / (2 * h)
        S = dT @ np.linalg.inv(T0)   # spatial twist matrix
        B = np.linalg.inv(T0) @ dT   # body twist matrix
        for M, J in ((S, Jw), (B, Jl)):
            w = np.stack([M[:, 2, 1], M[:, 0, 2], M[:, 1, 0]], axis=1)
            v = M[:, :3, 3]
            assert np.abs(J[:, :3, j] - v).max() < 1e-7 and np.abs(J[:, 3:, j] - w).max() < 1e-7
