"""CPU tests of the oracle's triangle routines and BVH distance: the SAT triangle-triangle intersection (FCL / PQP
TriContact) against an independent closed-form distance (vertex-face + edge-edge), including coplanar overlaps; the GJK
leaf distance against the same closed form; mesh / octree object distances against brute-force point sampling."""
import numpy as np

from tests.common import compile_desc, make_panda_world, random_q


def _rand_tris(rng, n, spread):
    A = rng.normal(size=(n, 3, 3)) * 0.3
    B = rng.normal(size=(n, 3, 3)) * 0.3 + rng.normal(size=(n, 1, 3)) * spread
    return A, B


def test_tri_tri_intersection_against_closed_form_distance():
    from oracle import oracle as O

    rng = np.random.default_rng(5)
    n_hit = n_miss = 0
    for spread in (0.1, 0.4, 0.8):
        A, B = _rand_tris(rng, 1500, spread)
        for a, b in zip(A, B):
            hit = O.tri_tri_intersect(a, b)
            ind = O.tri_tri_intersect(a, b, independent=True)
            assert hit == ind, (a, b)
            n_hit += hit
            n_miss += not hit
    assert n_hit > 300 and n_miss > 300
    # coplanar cases (the edge-piercing formulation of round 1 missed these; FCL's SAT does not)
    t = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0]], dtype=float)
    inside = np.array([[0.1, 0.1, 0], [0.4, 0.1, 0], [0.1, 0.4, 0]], dtype=float)
    crossing = np.array([[0.5, -0.5, 0], [0.5, 1.5, 0], [2.0, 0.5, 0]], dtype=float)
    apart = t + np.array([3.0, 0, 0])
    touching = t + np.array([1.0, 0, 0])  # shares the vertex (1, 0, 0)
    assert O.tri_tri_intersect(t, inside) and O.tri_tri_intersect(t, crossing) and O.tri_tri_intersect(t, touching)
    assert not O.tri_tri_intersect(t, apart)
    for other in (inside, crossing, touching):
        assert O.tri_tri_intersect(t, other, independent=True)
    assert O.tri_tri_distance(t, inside) == 0.0 and abs(O.tri_tri_distance(t, apart) - 2.0) < 1e-12
    # parallel planes, overlapping footprints
    assert not O.tri_tri_intersect(t, inside + np.array([0, 0, 0.01]))
    assert abs(O.tri_tri_distance(t, inside + np.array([0, 0, 0.01])) - 0.01) < 1e-12


def test_gjk_leaf_distance_against_closed_form():
    from oracle import oracle as O

    rng = np.random.default_rng(6)
    A, B = _rand_tris(rng, 3000, 0.8)
    n = 0
    for a, b in zip(A, B):
        d = O.tri_tri_distance(a, b, closed=True)  # valid for disjoint triangles
        g = O.tri_tri_distance(a, b, closed=False)
        if not O.tri_tri_intersect(a, b, independent=True):
            assert abs(g - d) < 1e-7, (d, g)
            n += 1
        else:
            assert g < 0
    assert n > 1000


def _sample_tri(rng, tri, k):
    u = rng.random((k, 2))
    m = u.sum(1) > 1
    u[m] = 1 - u[m]
    return tri[0] + u[:, :1] * (tri[1] - tri[0]) + u[:, 1:] * (tri[2] - tri[0])


def test_mesh_and_octree_object_distance_against_sampling():
    """distance(sphere, mesh) and distance(sphere, octree) from the exhaustive leaf minimum vs the distance to a dense
    point sampling of the same surfaces (an upper bound that converges from above)."""
    import os

    from mplib_b200.pymp import fcl
    from oracle import oracle as O
    from tests.common import ROOT
    from tests.workloads import box_surface_points

    w = make_panda_world()
    m = fcl.load_mesh_as_BVH(os.path.join(ROOT, "data/panda/franka_description/meshes/collision/link0.stl"))
    w.add_normal_object("mesh_obstacle", fcl.CollisionObject(m, [0.6, 0.3, 0.2], [1, 0, 0, 0]))
    rng = np.random.Generator(np.random.PCG64(4))
    pts = box_surface_points(rng, 4000, (0.4, 0.4, 0.05), (0.5, -0.3, 0.1))
    w.add_point_cloud("cloud", pts, 0.02)
    w.add_normal_object("probe", fcl.CollisionObject(fcl.Sphere(0.03), [0.2, 0.0, 0.9], [1, 0, 0, 0]))
    desc = compile_desc(w)
    names = desc.object_names
    i_mesh, i_cloud, i_probe = names.index("mesh_obstacle"), names.index("cloud"), names.index("probe")
    # a world with exactly the pairs probe x mesh, probe x cloud (fixed x fixed) -- poses varied through the description
    tri_world = (m.vertices[m.tris] + np.array([0.6, 0.3, 0.2]))
    cells = np.asarray(w.normal_objects["cloud"].geometry.cells, dtype=np.int64)  # octomap stores float32 points: use the tree's own voxel set
    rs = np.random.default_rng(1)
    surf = np.concatenate([_sample_tri(rs, t, 400) for t in tri_world])
    for trial in range(12):
        c = rs.uniform([0.2, -0.6, 0.0], [1.0, 0.7, 0.6])
        d2 = desc.with_objects([i_probe], obj_local=np.array([[1, 0, 0, c[0], 0, 1, 0, c[1], 0, 0, 1, c[2]]], dtype=float))
        d2 = d2.with_pairs([i_probe, i_probe], [i_mesh, i_cloud], ["x", "x"])
        md, am, pd = O.distance_batch(d2, np.zeros((1, d2.state_dim)), return_pairs=True)
        # mesh: sampled upper bound
        ub = np.linalg.norm(surf - c, axis=1).min() - 0.03
        if pd[0, 0] > 0:
            assert pd[0, 0] <= ub + 1e-9 and pd[0, 0] >= ub - 0.01, (pd[0, 0], ub)
        else:
            assert ub < 0.01
        # octree: exact point-to-box distance over the occupied voxels
        lo, hi = cells * 0.02, (cells + 1) * 0.02
        dd = np.linalg.norm(np.maximum(np.maximum(lo - c, c - hi), 0.0), axis=1).min() - 0.03
        if dd > 1e-9:
            assert abs(pd[0, 1] - dd) < 1e-7, (pd[0, 1], dd)
        else:
            assert pd[0, 1] == -1.0
