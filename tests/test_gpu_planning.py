"""GPU tests of the planning layer: batched motion validation vs the oracle's DiscreteMotionValidator restatement, and
lock-step RRTConnect (paths re-validated by the oracle at OMPL's resolution)."""
import numpy as np
import pytest

from tests.common import PANDA_HI, PANDA_LO, make_panda_world, random_q

pytestmark = pytest.mark.gpu


def _edges(n, seed, max_step):
    rng = np.random.Generator(np.random.PCG64(seed))
    a = random_q(n, seed)
    d = rng.uniform(-1, 1, a.shape) * rng.uniform(0, max_step, (n, 1))
    b = np.clip(a + d, PANDA_LO, PANDA_HI)
    return a, b


def _motion_parity(world, n, seed, max_step, max_mismatch):
    from mplib_b200.pymp import OMPLPlanner
    from oracle import oracle as O

    pl = OMPLPlanner(world)
    a, b = _edges(n, seed, max_step)
    a32, b32 = a.astype(np.float32), b.astype(np.float32)
    dev = world._sync()["dev"]
    valid, st = dev.check_motion(pl.space, a32, b32, stats=True)
    ov, ns = O.check_motion_batch(dev.desc, pl.lower_joint_limits, pl.upper_joint_limits, a32.astype(np.float64), b32.astype(np.float64),
                                  threads=8)
    mism = np.nonzero(valid != ov)[0]
    print(f"edges={n} valid-rate={ov.mean():.3f} gpu states={st['n_states']} mismatches={len(mism)}")
    # an edge may differ only through a sample inside the 1e-4 m contact band (or an fp32 ceil() at a segment-count
    # boundary moving the samples); both are rare
    assert len(mism) <= max_mismatch, mism[:10]
    # the device samples every interior point (no early exit): at least as many states as the oracle tested
    assert st["n_states"] >= ns * 0.5
    return valid


def test_motion_parity_self(panda_world):
    v = _motion_parity(panda_world, 20000, 41, 0.6, 6)
    assert 0.5 < v.mean() < 1.0


def test_motion_parity_c4_scene():
    from tests.workloads import c4_world

    _motion_parity(c4_world(cloud_points=20000), 4000, 42, 0.4, 4)


def test_motion_degenerate_edges(panda_world):
    from mplib_b200.pymp import OMPLPlanner

    pl = OMPLPlanner(panda_world)
    dev = panda_world._sync()["dev"]
    q = random_q(64, 43)
    coll = panda_world.collide_batch(q)
    v = dev.check_motion(pl.space, q, q)  # a == b: nd = 0, only b is tested
    assert np.array_equal(v, ~coll)
    assert dev.check_motion(pl.space, np.zeros((0, 7)), np.zeros((0, 7))).shape == (0,)
    with pytest.raises(RuntimeError):
        dev.check_motion(pl.space, np.zeros((3, 6)), np.zeros((3, 6)))


def _validate_paths(desc, pl, starts, goals, status, paths, range_):
    from oracle import oracle as O

    n_solved = 0
    for i, (s, p) in enumerate(zip(status, paths)):
        if s != "Exact solution":
            assert p.shape[0] == 0
            continue
        n_solved += 1
        assert p.shape[0] >= 2 and p.shape[1] == pl.dim
        assert np.allclose(p[0], starts[i].astype(np.float32), atol=1e-6) and np.allclose(p[-1], goals[i].astype(np.float32), atol=1e-6)
        step = np.abs(np.diff(p, axis=0)).sum(axis=1)
        assert step.max() <= range_ * (1 + 1e-4) + 1e-6, step.max()
        ok, _ = O.check_motion_batch(desc, pl.lower_joint_limits, pl.upper_joint_limits, p[:-1], p[1:], threads=8)
        # fp32 device vs double oracle: an edge can flip only inside the contact band; count, do not hide
        assert (~ok).sum() <= 1, (i, np.nonzero(~ok)[0])
    return n_solved


def test_rrtconnect_self_scene(panda_world):
    from mplib_b200.pymp import OMPLPlanner
    from tests.workloads import free_states

    pl = OMPLPlanner(panda_world)
    pts = free_states(panda_world, 128, 51, panda_world.collide_batch)
    starts, goals = pts[:64], pts[64:]
    status, paths, stats = pl.plan_batch(starts, goals, range=0.1, time=20.0, seed=7)
    print(stats, [len(p) for p in paths][:8])
    n = _validate_paths(panda_world._sync()["dev"].desc, pl, starts, goals, status, paths, 0.1)
    assert n == 64
    # deterministic for a fixed seed
    status2, paths2, _ = pl.plan_batch(starts, goals, range=0.1, time=20.0, seed=7)
    assert status2 == status and all(np.array_equal(a, b) for a, b in zip(paths, paths2))


def test_rrtconnect_c4_scene_and_plan_api():
    from mplib_b200.pymp import OMPLPlanner
    from tests.workloads import c4_world, free_states

    w = c4_world(cloud_points=20000)
    pl = OMPLPlanner(w)
    pts = free_states(w, 64, 52, w.collide_batch)
    starts, goals = pts[:32], pts[32:]
    status, paths, stats = pl.plan_batch(starts, goals, range=0.1, time=30.0, seed=3, max_rounds=20000)
    print(stats, status.count("Exact solution"))
    n = _validate_paths(w._sync()["dev"].desc, pl, starts, goals, status, paths, 0.1)
    assert n >= 28
    # single-query reference signature: (status, path[n, dim])
    st, path = pl.plan(starts[0], [goals[0]], range=0.1, time=10.0)
    assert st == "Exact solution" and path.shape[1] == 7
    # a colliding goal is never admitted to the goal tree -> "Invalid goal" (pis_.nextGoal() skips invalid states)
    bad = random_q(2000, 53)
    bad = bad[w.collide_batch(bad)][0]
    st, path = pl.plan(starts[0], [bad], range=0.1, time=1.0)
    assert st == "Invalid goal" and path.shape == (0, 7)
    with pytest.raises(RuntimeError):
        pl.plan(starts[0][:6], [goals[0][:6]])
    with pytest.raises(RuntimeError, match="Planner Not implemented"):
        pl.plan(starts[0], [goals[0]], planner_name="Nope")


def test_rrtconnect_slots_retire_and_refill():
    """Q queries through fewer concurrent slots: a slot whose query ends takes the next pending one.  The per-query
    result does not depend on the slot schedule (random numbers are keyed by the query, trees are per query), and the
    per-query round budget turns long searches into "Timeout" without touching the others."""
    from mplib_b200.pymp import OMPLPlanner
    from tests.workloads import c4_world, free_states

    w = c4_world(cloud_points=20000)
    pl = OMPLPlanner(w)
    pts = free_states(w, 192, 54, w.collide_batch)
    starts, goals = pts[:96], pts[96:]
    kw = dict(range=0.1, time=60.0, seed=11, max_rounds=200000, tree_capacity=4096)
    s_all, p_all, st_all = pl.plan_batch(starts, goals, **kw)
    s_16, p_16, st_16 = pl.plan_batch(starts, goals, n_slots=16, **kw)
    assert s_all == s_16
    assert all(np.array_equal(a, b) for a, b in zip(p_all, p_16))
    assert st_16["n_solved"] == st_all["n_solved"] >= 85 and st_16["rounds"] > st_all["rounds"]
    n = _validate_paths(w._sync()["dev"].desc, pl, starts, goals, s_16, p_16, 0.1)
    assert n == st_16["n_solved"]
    # per-query budget: the queries that needed more rounds end as "Timeout", the others are unchanged
    s_b, p_b, st_b = pl.plan_batch(starts, goals, n_slots=16, max_rounds_per_query=40, **kw)
    assert 0 < st_b["n_solved"] < st_all["n_solved"]
    for i, s in enumerate(s_b):
        if s == "Exact solution":
            assert s_all[i] == s and np.array_equal(p_b[i], p_all[i])
        else:
            assert s == "Timeout" and p_b[i].shape == (0, 7)
