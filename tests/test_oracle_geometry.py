"""CPU tests pinning the oracle's narrowphase: analytic primitive cases, brute-force SAT
as a second opinion on the MPR / GJK restatements on the Panda hulls (8c (4)), octomap voxel rule (8c (6))."""
import numpy as np
import pytest

from mplib_b200.scene import (GEOM_BOX, GEOM_CAPSULE, GEOM_CONVEX, GEOM_CYLINDER, GEOM_SPHERE, SceneBuilder)
from tests.common import compile_desc, random_q


def _T(p=(0, 0, 0), R=None):
    T = np.eye(4)
    T[:3, 3] = p
    if R is not None:
        T[:3, :3] = R
    return T


def _rot(axis, ang):
    axis = np.asarray(axis, dtype=float)
    axis /= np.linalg.norm(axis)
    K = np.array([[0, -axis[2], axis[1]], [axis[2], 0, -axis[0]], [-axis[1], axis[0], 0]])
    return np.eye(3) + np.sin(ang) * K + (1 - np.cos(ang)) * K @ K


@pytest.fixture(scope="module")
def prim_desc():
    b = SceneBuilder()
    ids = {}
    ids["s1"] = b.add_object(GEOM_SPHERE, -1, np.eye(4), (0.1,), name="s1")
    ids["s2"] = b.add_object(GEOM_SPHERE, -1, np.eye(4), (0.25,), name="s2")
    ids["b1"] = b.add_object(GEOM_BOX, -1, np.eye(4), (0.2, 0.4, 0.6), name="b1")
    ids["b2"] = b.add_object(GEOM_BOX, -1, np.eye(4), (0.1, 0.1, 0.1), name="b2")
    ids["c1"] = b.add_object(GEOM_CAPSULE, -1, np.eye(4), (0.05, 0.3), name="c1")
    ids["c2"] = b.add_object(GEOM_CAPSULE, -1, np.eye(4), (0.02, 0.1), name="c2")
    ids["y1"] = b.add_object(GEOM_CYLINDER, -1, np.eye(4), (0.05, 0.3), name="y1")
    # unit cube as a convex hull, to compare hull code with the box closed forms
    v = np.array([[x, y, z] for x in (-.5, .5) for y in (-.5, .5) for z in (-.5, .5)]) * [0.2, 0.4, 0.6]
    f = [[0, 1, 3], [0, 3, 2], [4, 6, 7], [4, 7, 5], [0, 4, 5], [0, 5, 1], [2, 3, 7], [2, 7, 6], [0, 2, 6], [0, 6, 4],
         [1, 5, 7], [1, 7, 3]]
    ids["h1"] = b.add_object(GEOM_CONVEX, -1, np.eye(4), geom=b.add_hull(v, f), name="h1")
    return b.finish(), ids


def test_sphere_sphere(prim_desc):
    from oracle import oracle as O

    d, i = prim_desc
    for gap in (-0.05, -1e-6, 1e-6, 0.05, 0.7):
        Tb = _T((0.35 + gap, 0, 0))
        assert O.object_collide(d, i["s1"], _T(), i["s2"], Tb) == (gap <= 0)
        if gap > 0:
            assert abs(O.object_distance(d, i["s1"], _T(), i["s2"], Tb) - gap) < 2e-6


def test_sphere_box_face_edge_corner(prim_desc):
    from oracle import oracle as O

    d, i = prim_desc
    h = np.array([0.1, 0.2, 0.3])
    R = _rot([1, 2, 3], 0.7)
    Tb = _T((0.3, -0.2, 0.1), R)
    for direction in ([1, 0, 0], [1, 1, 0], [1, 1, 1], [0, -1, 1]):
        n = np.asarray(direction, dtype=float)
        corner = h * np.sign(n)  # closest feature of the box along that direction family
        n /= np.linalg.norm(n)
        for gap in (-0.02, 1e-5, 0.03):
            c_local = corner + n * (0.1 + gap)
            Ts = _T(Tb[:3, :3] @ c_local + Tb[:3, 3])
            assert O.object_collide(d, i["s1"], Ts, i["b1"], Tb) == (gap <= 0)
            assert O.object_collide(d, i["b1"], Tb, i["s1"], Ts) == (gap <= 0)
            if gap > 0:
                assert abs(O.object_distance(d, i["s1"], Ts, i["b1"], Tb) - gap) < 2e-6
    # sphere centre inside the box
    assert O.object_collide(d, i["s1"], _T(Tb[:3, 3]), i["b1"], Tb)


def test_box_box_axis_aligned_and_rotated(prim_desc):
    from oracle import oracle as O

    d, i = prim_desc
    # axis aligned along x: half extents 0.1 and 0.05
    for gap in (-0.01, -1e-7, 1e-7, 0.01):
        Tb = _T((0.15 + gap, 0.03, -0.1))
        assert O.object_collide(d, i["b1"], _T(), i["b2"], Tb) == (gap <= 0)
        assert O.object_collide(d, i["h1"], _T(), i["b2"], Tb) == (gap <= 0)  # hull (MPR) agrees with SAT
    # b2 rotated 45 deg about z: its extent along x is 0.05*sqrt(2)
    R = _rot([0, 0, 1], np.pi / 4)
    ext = 0.05 * np.sqrt(2)
    for gap in (-0.01, 1e-6, 0.01):
        Tb = _T((0.1 + ext + gap, 0, 0), R)
        assert O.object_collide(d, i["b1"], _T(), i["b2"], Tb) == (gap <= 0)
        if gap > 0:
            assert abs(O.object_distance(d, i["b1"], _T(), i["b2"], Tb) - gap) < 2e-6
    # edge-edge separation: needs a cross-product axis.  b2 spun about x by 45 deg then placed diagonally in y-z
    # b1's edge along x at (y, z) = (0.2, 0.3) against an edge of b2 along e2 = (0, 1, -1)/sqrt2: the separating
    # axis is x cross e2 = n = (0, 1, 1)/sqrt2, none of the 6 face normals separates them
    n = np.array([0, 1, 1]) / np.sqrt(2)
    x = np.array([1.0, 0, 0])
    u, v, w = np.array([0, 1, -1]) / np.sqrt(2), (n + x) / np.sqrt(2), (n - x) / np.sqrt(2)
    R = np.stack([v, w, u], axis=1)
    if np.linalg.det(R) < 0:
        R = np.stack([w, v, u], axis=1)
    edge_pt = np.array([0, 0.2, 0.3])
    for gap in (-0.01, -1e-5, 1e-5, 0.01):
        Tb = _T(edge_pt + n * (0.05 * np.sqrt(2) + gap), R)
        assert O.object_collide(d, i["b1"], _T(), i["b2"], Tb) == (gap <= 0)
        assert O.object_collide(d, i["h1"], _T(), i["b2"], Tb) == (gap <= 0)
        if gap > 0:
            assert abs(O.object_distance(d, i["b1"], _T(), i["b2"], Tb) - gap) < 2e-6


def test_capsule_capsule_and_cylinder(prim_desc):
    from oracle import oracle as O

    d, i = prim_desc
    # parallel capsules side by side
    for gap in (-0.01, 1e-5, 0.02):
        Tb = _T((0.07 + gap, 0, 0.02))
        assert O.object_collide(d, i["c1"], _T(), i["c2"], Tb) == (gap <= 0)
        if gap > 0:
            assert abs(O.object_distance(d, i["c1"], _T(), i["c2"], Tb) - gap) < 2e-6
    # crossed capsules (perpendicular axes): distance = offset - r1 - r2
    R = _rot([1, 0, 0], np.pi / 2)
    for gap in (-0.01, 0.02):
        Tb = _T((0.07 + gap, 0, 0), R)
        assert O.object_collide(d, i["c1"], _T(), i["c2"], Tb) == (gap <= 0)
    # end to end: hemispherical caps
    for gap in (-0.005, 0.01):
        Tb = _T((0, 0, 0.15 + 0.05 + 0.05 + 0.02 + gap))
        assert O.object_collide(d, i["c1"], _T(), i["c2"], Tb) == (gap <= 0)
        if gap > 0:
            assert abs(O.object_distance(d, i["c1"], _T(), i["c2"], Tb) - gap) < 2e-6
    # sphere vs cylinder: flat cap, side, rim
    for pos, gapdir in (((0, 0, 0.15 + 0.1), (0, 0, 1)), ((0.05 + 0.1, 0, 0.05), (1, 0, 0))):
        for gap in (-0.01, 0.01):
            Ts = _T(np.asarray(pos) + gap * np.asarray(gapdir, dtype=float))
            assert O.object_collide(d, i["s1"], Ts, i["y1"], _T()) == (gap <= 0)
    rim = np.array([0.05, 0, 0.15])
    n = np.array([1, 0, 1]) / np.sqrt(2)
    for gap in (-0.01, 0.01):
        assert O.object_collide(d, i["s1"], _T(rim + n * (0.1 + gap)), i["y1"], _T()) == (gap <= 0)


def test_sat_second_opinion_on_panda_hulls(panda_world):
    """>= 1e4 random pair poses.  MPR verdict vs exhaustive SAT (face normals + edge x edge):
    they may differ only inside the 1e-4 m contact band."""
    from oracle import oracle as O

    desc = compile_desc(panda_world)
    q = random_q(2000, 31)
    v, ph = O.collide_batch(desc, q, threads=4)
    # SAT is O(E^2) per pair: run it on every colliding pair plus a stride of separated ones
    rng = np.random.Generator(np.random.PCG64(9))
    take = ph | (rng.random(ph.shape) <= 0.28)
    items = np.argwhere(take)
    gaps = O.sat_batch(desc, q, items[:, 0], items[:, 1], threads=8)
    checked, near = len(items), int((np.abs(gaps) < 1e-4).sum())
    for (i, p), gap in zip(items, gaps):
        if abs(gap) >= 1e-4:
            assert (gap < 0) == bool(ph[i, p]), (i, p, gap, ph[i, p])
    assert checked >= 10000
    print(f"SAT second opinion: {checked} pair poses, {near} inside the contact band")


def test_gjk_distance_vs_sat_lower_bound(panda_world):
    """Separated hull pairs: GJK distance >= SAT gap (SAT gives a lower bound that is tight for face contacts)."""
    from oracle import oracle as O

    desc = compile_desc(panda_world)
    q = random_q(60, 32)
    for i in range(q.shape[0]):
        for p in range(0, desc.n_pairs, 3):
            sd = O.pair_signed_distance(desc, q[i], p)
            gap = O.sat_pair(desc, q[i], p)
            if sd > 1e-4:
                assert gap > 0 and sd >= gap - 1e-6
            elif sd < -1e-4:
                assert gap < 0


def test_octomap_cells_rule():
    """every touched finest-level voxel floor(p/res) is occupied; duplicates collapse;
    float32 point storage; out-of-range keys are dropped."""
    from mplib_b200.pymp import fcl

    pts = np.array([[0.0049, 0.0149, -0.0001], [0.0051, 0.0101, -0.0099], [0.0151, 0.0, 0.0], [0.0149, 0.0, 0.0],
                    [400.0, 0, 0]])
    cells = fcl.octomap_cells(pts, 0.01)
    assert sorted(map(tuple, cells)) == sorted({(0, 1, -1), (1, 0, 0)})
    rng = np.random.Generator(np.random.PCG64(3))
    p = rng.uniform(-1, 1, (5000, 3))
    res = 0.013
    got = set(map(tuple, fcl.OcTree(p, res).cells))
    want = set(map(tuple, np.floor(p.astype(np.float32).astype(np.float64) / res).astype(int)))
    assert got == want


def test_octree_oracle_matches_exact_voxel_test():
    """Oracle octree x shape vs an independent numpy box-sphere test over all voxels."""
    from oracle import oracle as O
    from mplib_b200.scene import GEOM_OCTREE

    rng = np.random.Generator(np.random.PCG64(5))
    pts = rng.uniform(-0.2, 0.2, (300, 3))
    res = 0.02
    cells = np.unique(np.floor(pts.astype(np.float32).astype(np.float64) / res).astype(np.int32), axis=0)
    b = SceneBuilder()
    oc = b.add_object(GEOM_OCTREE, -1, np.eye(4), geom=b.add_octree(cells, res), name="oct")
    sp = b.add_object(GEOM_SPHERE, -1, np.eye(4), (0.03,), name="s")
    d = b.finish()
    cen = (cells + 0.5) * res
    To = _T((0.05, -0.02, 0.01), _rot([0.3, 1, -0.5], 1.1))
    n_hit = 0
    for k in range(400):
        c = rng.uniform(-0.3, 0.3, 3)
        local = To[:3, :3].T @ (c - To[:3, 3])
        dist = np.linalg.norm(np.maximum(np.abs(local - cen) - res / 2, 0), axis=1).min()
        if abs(dist - 0.03) < 1e-6:
            continue
        assert O.object_collide(d, oc, To, sp, _T(c)) == (dist < 0.03)
        assert O.object_collide(d, sp, _T(c), oc, To) == (dist < 0.03)
        n_hit += dist < 0.03
    assert 20 < n_hit < 380
