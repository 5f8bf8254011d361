"""GPU distance path (mpb_distance_batch) vs the oracle's GJK restatement: distances within 1e-4 m, -1 verdicts equal
outside the contact band, arg-min pair consistent."""
import numpy as np
import pytest

from tests.common import make_panda_world, random_q

pytestmark = pytest.mark.gpu
TOL = 1e-4


def _distance_parity(world, q):
    from oracle import oracle as O

    c = world._sync()
    dev, desc = c["dev"], c["dev"].desc
    md, am, st = dev.distance(q, stats=True)
    qd = q.astype(np.float32).astype(np.float64)
    omd, oam = O.distance_batch(desc, qd)
    n_band = 0
    for i in range(q.shape[0]):
        g, o = md[i], omd[i]
        if (g < 0) != (o < 0):  # collision verdict differs: only allowed inside the contact band
            other = o if g < 0 else g
            assert other < TOL, (i, g, o)
            n_band += 1
            continue
        if g < 0:
            continue
        assert abs(g - o) < TOL, (i, g, o, am[i], oam[i])
        if am[i] != oam[i]:  # two pairs nearly tie: the GPU's pair must be (nearly) as close in the oracle too
            sd = O.pair_signed_distance(desc, qd[i], int(am[i]))
            assert abs(sd - o) < 2 * TOL, (i, am[i], oam[i], sd, o)
    print(f"n={q.shape[0]} collide-rate={(omd < 0).mean():.3f} band={n_band} median={np.median(omd[omd > 0]):.4f} stats={st}")
    # items that stop at the iteration cap (fp32 cycling between two simplices) still passed the 1e-4 check above
    # (cylinder rims converge only asymptotically); they are counted, not hidden
    assert st["narrow_capped"] <= 3 + st["pairs_after_cull"] // 500
    return md, am


def test_distance_self(panda_world):
    md, am = _distance_parity(panda_world, random_q(4000, 71))
    # -1 exactly where collide() says so (same scene, same admitted pairs), up to the contact band
    v = panda_world.collide_batch(random_q(4000, 71))
    assert ((md < 0) != v).sum() <= 2


def test_distance_two_pandas():
    from tests.workloads import c5

    wl = c5(1024)
    _distance_parity(wl["world"], wl["make_q"](0, 1024))


def test_distance_primitives_and_api():
    from mplib_b200.pymp import fcl
    from tests.workloads import add_c3_primitives, attach_c3_box

    w = make_panda_world()
    add_c3_primitives(w, 4, seed=3)  # 4 spheres, 4 boxes, 4 capsules, 4 cylinders
    attach_c3_box(w)
    _distance_parity(w, random_q(1500, 72))
    # legacy single-state API (python/pybind_planning_world.hpp:117-123)
    w.set_qpos_all([0, 0.19, 0.0, -2.61, 0.0, 2.94, 0.78])
    r = w.distance_full()
    assert r.min_distance == w.distance() and r.distance_type != ""
    rs, ro = w.self_distance(), w.distance_with_others()
    assert rs.distance_type in ("self", "self_attach") and ro.distance_type in ("articulation_sceneobject", "attach_sceneobject")
    assert abs(min(rs.min_distance, ro.min_distance) - r.min_distance) < 1e-6
    # the pair list is restored after a sectioned query
    assert w._sync()["dev"].n_pairs == len(w._sync()["pairs"])
    # meshes / octrees are refused loudly, not approximated
    w.add_point_cloud("cloud", np.random.default_rng(0).uniform(-1, 1, (100, 3)), 0.05)
    with pytest.raises(Exception, match="not supported"):
        w.distance()
