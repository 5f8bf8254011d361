"""GPU parity: CUDA path (through the C-ABI) vs the CPU oracle on the same seeded inputs."""
import numpy as np
import pytest

from tests.common import compile_desc, make_panda_world, random_q

pytestmark = pytest.mark.gpu

BAND = 1e-4  # north_star: verdicts may differ only where |signed distance| < 1e-4 m


def _check_pairs(desc, q, gpu_pairs, orc_pairs):
    from oracle import oracle as O

    mism = np.argwhere(gpu_pairs != orc_pairs)
    in_band = 0
    for i, p in mism:
        sd = O.pair_signed_distance(desc, q[i], int(p))
        if abs(sd) >= BAND:
            sat = O.sat_pair(desc, q[i], int(p))
            assert abs(sat) < BAND, f"state {i} pair {p}: gpu={gpu_pairs[i, p]} oracle={orc_pairs[i, p]} sd={sd} sat={sat}"
        in_band += 1
    return len(mism), in_band


def test_fk_parity(panda_world):
    from oracle import oracle as O

    q = random_q(4096, 11)
    c = panda_world._sync()
    got = c["dev"].fk(q)
    ref = O.fk_batch(c["dev"].desc, q.astype(np.float32).astype(np.float64))
    assert np.abs(got[..., :3] - ref[..., :3]).max() < 1e-5
    # quaternion sign is not unique near the branch boundaries: compare up to sign
    dq = np.minimum(np.abs(got[..., 3:] - ref[..., 3:]).max(-1), np.abs(got[..., 3:] + ref[..., 3:]).max(-1))
    assert dq.max() < 1e-5


def test_self_collision_parity(panda_world):
    from oracle import oracle as O

    q = random_q(20000, 1)
    c = panda_world._sync()
    desc = c["dev"].desc
    v, ph, st = c["dev"].collide(q, want_pairs=True, stats=True)
    qd = q.astype(np.float32).astype(np.float64)
    ov, oph = O.collide_batch(desc, qd)
    n_mis, n_band = _check_pairs(desc, qd, ph, oph)
    print(f"mismatches={n_mis} (all in band), collision rate={ov.mean():.3f}, stats={st}")
    assert n_mis <= 20
    assert st["narrow_capped"] == 0
    # verdict bit == OR over pair bits
    assert np.array_equal(v, ph.any(axis=1))


def test_example_poses(panda_world):
    """examples/detect_collision.py:22-34"""
    w = panda_world
    w.set_qpos_all([0, 0.19, 0.0, -2.61, 0.0, 2.94, 0.78])
    assert w.self_collide() == [] and not w.collide()
    w.set_qpos_all([0, 1.36, 0, -3, -3, 3, -1])
    r = w.self_collide()
    assert len(r) > 0 and all(x.collision_type == "self" for x in r)
    assert w.collide()


def test_ragged_and_empty(panda_world):
    dev = panda_world._sync()["dev"]
    v, ph = dev.collide(np.zeros((0, 7)))
    assert v.shape == (0,)
    for n in (1, 31, 33, 1000):
        q = random_q(n, 5)
        v, ph = dev.collide(q)
        v2, ph2 = dev.collide(q[::-1].copy())
        assert np.array_equal(v, v2[::-1]) and np.array_equal(ph, ph2[::-1])


def _env_parity(world, n, seed, max_mismatch):
    _env_parity_q(world, random_q(n, seed), max_mismatch)


def _env_parity_q(world, q, max_mismatch):
    from oracle import oracle as O

    n = q.shape[0]
    c = world._sync()
    desc = c["dev"].desc
    v, ph, st = c["dev"].collide(q, want_pairs=True, stats=True)
    qd = q.astype(np.float32).astype(np.float64)
    ov, oph = O.collide_batch(desc, qd, threads=8)
    mism = np.argwhere(ph != oph)
    bad = []
    for i, p in mism:
        sd = O.pair_signed_distance(desc, qd[i], int(p))
        if not (abs(sd) < BAND):
            bad.append((int(i), int(p), desc.pair_kinds[p], desc.object_names[desc.pair_a[p]], desc.object_names[desc.pair_b[p]],
                        bool(ph[i, p]), bool(oph[i, p]), sd))
    print(f"pairs={desc.n_pairs} mismatches={len(mism)} out-of-band={len(bad)} collide-rate={ov.mean():.3f} stats={st}")
    assert not bad, bad[:10]
    assert len(mism) <= max_mismatch
    assert np.array_equal(v, ph.any(axis=1))
    assert st["narrow_capped"] <= max(2, n // 1000)
    # verdict-only mode (early exits enabled) must give the same verdicts
    v2, _ = c["dev"].collide(q, want_pairs=False)
    assert np.array_equal(v2, v)


def test_env_primitives_attached():
    """C3 minus mesh and cloud: 128 primitives + grasped box (AttachedBody), ACM touch links."""
    from tests.workloads import c3_world

    _env_parity(c3_world(mesh=False, cloud_points=0), 3000, 21, 30)


def test_env_octree():
    from tests.workloads import c3_world

    _env_parity(c3_world(primitives=False, mesh=False, cloud_points=20000, attach=True), 3000, 22, 30)


def test_env_mesh():
    from tests.workloads import c3_world

    _env_parity(c3_world(primitives=False, mesh=True, cloud_points=0, attach=True), 3000, 23, 30)


def test_env_full_c3():
    from tests.workloads import c3_world

    _env_parity(c3_world(), 1500, 24, 30)


def test_two_panda_c5():
    """configs[4]: two planned articulations, name-ordered 14-dof state, cross-robot pairs with ACM entries."""
    from oracle import oracle as O
    from tests.workloads import c5

    wl = c5(4096)
    w, q = wl["world"], wl["make_q"](0, 4096)
    _env_parity_q(w, q, 30)
    # FK of both robots (26 link frames) against the oracle
    dev = w._sync()["dev"]
    got = dev.fk(q[:512])
    ref = O.fk_batch(dev.desc, q[:512].astype(np.float64))
    assert np.abs(got[..., :3] - ref[..., :3]).max() < 1e-5
    # an ACM edit re-uses the compiled scene with a new pair list
    n0 = dev.n_pairs
    w.get_allowed_collision_matrix().set_entry("a_panda_hand", "b_panda_hand", True)
    dev2 = w._sync()["dev"]
    assert dev2 is dev and dev2.n_pairs == n0 - 1
    _env_parity_q(w, q[:1024], 30)


def _golden_parity(world, fixture, max_mismatch):
    """CUDA path vs the committed oracle vectors of tests/golden/ (no oracle run except the signed distance of a
    mismatching pair, to prove it lies inside the contact band)."""
    from oracle import oracle as O
    from tests.test_golden_cpu import load_golden, pair_names

    q, v, ph, names = load_golden(fixture)
    c = world._sync()
    desc = c["dev"].desc
    assert names == pair_names(desc)
    gv, gph = c["dev"].collide(q, want_pairs=True)
    mism = np.argwhere(gph != ph)
    for i, p in mism:
        sd = O.pair_signed_distance(desc, q[i].astype(np.float64), int(p))
        assert abs(sd) < BAND, (fixture, int(i), names[p], sd)
    print(f"{fixture}: states={q.shape[0]} pairs={len(names)} mismatches={len(mism)} (all inside the 1e-4 m band)")
    assert len(mism) <= max_mismatch
    assert np.array_equal(gv, gph.any(axis=1))


def test_golden_c1_scene_a():
    """configs[0], scene A: examples/detect_collision.py floor point cloud (octree at 1e-3), 10k qpos."""
    from tests.workloads import c1_world

    _golden_parity(c1_world("A"), "c1_sceneA.npz", 20)


def test_golden_c1_scene_b():
    """configs[0], scene B: box + sphere obstacles, 10k qpos."""
    from tests.workloads import c1_world

    _golden_parity(c1_world("B"), "c1_sceneB.npz", 20)


def test_golden_c2_self(panda_world):
    _golden_parity(panda_world, "c2_self_20k.npz", 20)


def test_golden_c5_two_panda():
    from tests.workloads import c5_world

    _golden_parity(c5_world(), "c5_two_panda_4k.npz", 20)


# ---------------------------------------------------------------------------------------------- mesh links (convex=False)
_DIRS26 = np.array([[x, y, z] for x in (-1, 0, 1) for y in (-1, 0, 1) for z in (-1, 0, 1) if (x, y, z) != (0, 0, 0)], dtype=np.float64)
_DIRS26 /= np.linalg.norm(_DIRS26, axis=1, keepdims=True)


def _in_band_by_perturbation(desc, poses, a, b, step=1.13e-4):
    """Signed distance is 1-Lipschitz in a translation of one object, so a pair whose oracle verdict flips under
    a translation of `step` (26 directions; 1.13e-4 = 1e-4 / cos of the largest angle to the nearest direction) has
    |signed distance| < step.  Used where the oracle has no signed-distance routine (mesh x mesh, mesh x octree)."""
    from oracle import oracle as O

    base = O.object_collide(desc, a, poses[a], b, poses[b])
    for d in _DIRS26:
        Tb = poses[b].copy()
        Tb[:, 3] += step * d
        if O.object_collide(desc, a, poses[a], b, Tb) != base:
            return True
    return False


def _mesh_parity(world, q, max_mismatch):
    from oracle import oracle as O

    c = world._sync()
    desc = c["dev"].desc
    v, ph, st = c["dev"].collide(q, want_pairs=True, stats=True)
    qd = q.astype(np.float32).astype(np.float64)
    ov, oph = O.collide_batch(desc, qd, threads=8)
    mism = np.argwhere(ph != oph)
    bad = []
    for i, p in mism:
        poses = O.object_poses(desc, qd[i:i + 1])[0]
        if not _in_band_by_perturbation(desc, poses, int(desc.pair_a[p]), int(desc.pair_b[p])):
            bad.append((int(i), int(p), desc.object_names[desc.pair_a[p]], desc.object_names[desc.pair_b[p]], bool(ph[i, p]), bool(oph[i, p])))
    print(f"pairs={desc.n_pairs} mismatches={len(mism)} out-of-band={len(bad)} collide-rate={ov.mean():.3f} stats={st}")
    assert not bad, bad[:10]
    assert len(mism) <= max_mismatch
    assert np.array_equal(v, ph.any(axis=1))
    assert st["narrow_capped"] == 0
    v2, _ = c["dev"].collide(q, want_pairs=False)
    assert np.array_equal(v2, v)
    return ov


def test_mesh_links_self_collision():
    """BVHModel links (`convex=False`, tests/test_articulated_model.cpp:14-15): mesh x mesh self pairs."""
    w = make_panda_world(convex=False)
    ov = _mesh_parity(w, random_q(600, 31), 12)
    assert 0.01 < ov.mean() < 0.5


def test_mesh_links_vs_mesh_octree_and_primitives():
    """mesh links against a mesh obstacle (mesh x mesh), a point cloud (octree x mesh) and primitives."""
    import os

    from mplib_b200.pymp import fcl
    from tests.common import ROOT
    from tests.workloads import add_c3_primitives, box_surface_points

    w = make_panda_world(convex=False)
    m = fcl.load_mesh_as_BVH(os.path.join(ROOT, "data/panda/franka_description/meshes/collision/link0.stl"))
    w.add_normal_object("mesh_obstacle", fcl.CollisionObject(m, [0.45, 0.2, 0.3], [1, 0, 0, 0]))
    rng = np.random.Generator(np.random.PCG64(4))
    w.add_point_cloud("cloud", box_surface_points(rng, 4000, (0.8, 0.8, 0.05), (0.56, 0, -0.025)), 0.01)
    add_c3_primitives(w, 2)
    ov = _mesh_parity(w, random_q(300, 32), 12)
    assert ov.mean() > 0.05


def test_expand_variants_agree_and_match_oracle():
    """A large pass takes the lane-per-item expansion (expand_lane_kernel) and the thread-per-configuration sphere phase,
    small passes the warp-per-item expansion and the thread-per-pair sphere phase: bit-identical pair masks, and the
    large pass against the oracle."""
    from oracle import oracle as O
    from tests.workloads import c3_world

    w = c3_world()
    dev = w._sync()["dev"]
    q = random_q(24000, 31)
    v, ph, st = dev.collide(q, want_pairs=True, stats=True)
    for k in range(0, 24000, 8000):
        v2, ph2 = dev.collide(q[k:k + 2000], want_pairs=True)
        assert np.array_equal(ph[k:k + 2000], ph2) and np.array_equal(v[k:k + 2000], v2)
    qd = q[:6000].astype(np.float32).astype(np.float64)
    ov, oph = O.collide_batch(dev.desc, qd, threads=8)
    mism = np.argwhere(ph[:6000] != oph)
    for i, p in mism:
        assert abs(O.pair_signed_distance(dev.desc, qd[i], int(p))) < BAND
    assert len(mism) <= 30 and st["narrow_capped"] == 0


def _base_ring_world():
    """C3 primitives (grid scene: > 256 world-fixed partners) plus a ring of small spheres at the surface of the
    world-fixed base link: fixed x fixed pairs whose verdict must not depend on the broadphase grid."""
    from mplib_b200.pymp import fcl
    from tests.workloads import c3_world

    w = c3_world(n_each=32, mesh=False, cloud_points=0)
    k = 0
    for r in (0.07, 0.10, 0.13, 0.16):
        for az in np.linspace(0.0, 2.0 * np.pi, 8, endpoint=False):
            for z in (0.03, 0.09):
                pos = [-0.04 + r * np.cos(az), r * np.sin(az), z]
                w.add_normal_object(f"ring_{k}", fcl.CollisionObject(fcl.Sphere(0.03), pos, [1, 0, 0, 0]))
                k += 1
    return w


def test_world_fixed_link_against_obstacles_in_a_grid_scene():
    from oracle import oracle as O

    w = _base_ring_world()
    dev = w._sync()["dev"]
    desc = dev.desc
    q = random_q(6000, 41)  # > 4096: the thread-per-configuration kernels (sphere_grid_kernel)
    v, ph = dev.collide(q, want_pairs=True)
    qd = q[:64].astype(np.float32).astype(np.float64)
    ov, oph = O.collide_batch(desc, qd, threads=8)
    world_fixed = np.array([l < 0 or desc.link_joint[l] == 0 for l in desc.obj_link])  # obstacles, and links welded to the world
    fixed = world_fixed[np.asarray(desc.pair_a)] & world_fixed[np.asarray(desc.pair_b)]
    assert fixed.sum() >= 64  # base link x (primitives + ring)
    # fixed x fixed verdicts are the same for every configuration, and equal to the oracle's
    assert np.all(ph[:, fixed] == ph[0, fixed])
    assert np.array_equal(ph[0, fixed], oph[0, fixed])
    assert 4 <= int(ph[0, fixed].sum()) < int(fixed.sum())  # some of the ring touches the base, some does not
    n_mis, _ = _check_pairs(desc, qd, ph[:64], oph)
    assert n_mis <= 5
