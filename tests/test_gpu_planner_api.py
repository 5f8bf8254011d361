"""GPU tests of the Planner mirror (mplib/planner.py surface on the accelerated path): collision queries, SRDF
auto-generation statistic vs the oracle, joint-space planning."""
import os

import numpy as np
import pytest

from tests.common import PANDA_JOINTS, PANDA_LINKS, ROOT, SRDF, URDF, random_q

pytestmark = pytest.mark.gpu


def _planner(**kw):
    from mplib_b200.planner import Planner

    return Planner(URDF, PANDA_LINKS, PANDA_JOINTS, "panda_hand", np.ones(7), np.ones(7), srdf=SRDF, **kw)


def test_check_for_collision_calls():
    """examples/detect_collision.py:22-65 through Planner.check_for_self_collision / check_for_env_collision."""
    from tests.workloads import box_surface_points

    pl = _planner()
    free = [0, 0.19, 0.0, -2.61, 0.0, 2.94, 0.78, 0, 0]
    assert pl.check_for_self_collision(free) == []
    hits = pl.check_for_self_collision([0, 1.36, 0, -3, -3, 3, -1, 0, 0])
    assert hits and all(h.collision_type == "self" for h in hits)
    # 7-long qpos: finger joints are taken from the current state (mplib/planner.py:212-216)
    assert pl.check_for_self_collision(free[:7]) == []
    rng = np.random.Generator(np.random.PCG64(100))
    pl.update_point_cloud(box_surface_points(rng, 10000, (2.0, 2.0, 0.1), (0.0, 0.0, -0.1)), 1e-3)
    assert pl.check_for_env_collision(free) == []
    env = pl.check_for_env_collision([0, 1.5, 0, -1.5, 0, 0, 0, 0, 0])  # arm reaching below the floor
    assert env and all(h.collision_type == "articulation_sceneobject" for h in env)
    # the robot state is restored after every query
    assert np.allclose(pl.robot.get_qpos(), 0.0)


def test_pair_statistic_matches_oracle(tmp_path):
    """generate_collision_pair's statistic (mplib/planner.py:118-163) as one batched collide_full: per-pair collision
    counts over random configurations of a Panda WITHOUT an SRDF (46 pairs) equal the oracle's column sums."""
    from mplib_b200.planner import Planner
    from oracle import oracle as O

    d = str(tmp_path)
    os.symlink(os.path.join(ROOT, "data/panda/franka_description"), os.path.join(d, "franka_description"))
    urdf = os.path.join(d, "panda_nosrdf.urdf")
    with open(urdf, "w") as f:
        f.write(open(URDF).read())
    pl = Planner(urdf, PANDA_LINKS, PANDA_JOINTS, "panda_hand", np.ones(7), np.ones(7), srdf="", srdf_sample_time=20000)
    assert os.path.exists(pl.srdf)
    txt = open(pl.srdf).read()
    # the written SRDF holds the always-colliding pairs only; adjacent links are already removed by the URDF parent rule
    n_disabled = txt.count("<disable_collisions")
    # fresh planner on the generated SRDF: its pair list = 46 - n_disabled
    w = pl.planning_world
    assert len(w.get_pair_metadata()) == 46 - n_disabled
    # statistic vs oracle on the same states (full 9-dof states: all joints sampled)
    pl2 = Planner(urdf, PANDA_LINKS, PANDA_JOINTS, "panda_hand", np.ones(7), np.ones(7), srdf=pl.srdf)
    q = random_q(20000, 7)
    cnt = pl2.pair_collision_counts(q)
    desc = pl2.planning_world.get_scene_desc()
    _, oph = O.collide_batch(desc, q.astype(np.float32).astype(np.float64), threads=8)
    meta = pl2.planning_world.get_pair_metadata()
    ocnt = np.zeros_like(cnt)
    for p, c in enumerate(oph.sum(axis=0)):
        ocnt[pl2.link_name_2_idx[meta[p][4]], pl2.link_name_2_idx[meta[p][5]]] += int(c)
    assert np.abs(cnt - ocnt).sum() <= 20  # contact-band flips only
    # hand x link7 survives the URDF parent rule (the hand's link parent is the geometry-less link8) and always collides:
    # it is what the generated SRDF must disable (data/panda/panda.srdf lists it as "Adjacent")
    assert 'link1="panda_link7" link2="panda_hand"' in txt.replace("\n", " ")


def test_plan_qpos():
    from tests.workloads import free_states

    pl = _planner()
    pts = free_states(None, 2, 61, pl.check_for_collision_batch)
    start9 = np.concatenate([pts[0], [0, 0]])
    res = pl.plan_qpos([pts[1]], start9, planning_time=10.0)
    assert res["status"] == "Success"
    path = res["position"]
    assert path.shape[1] == 7 and np.allclose(path[0], pts[0], atol=1e-6) and np.allclose(path[-1], pts[1], atol=1e-6)
    assert not pl.check_for_collision_batch(path).any()
    out, stats = pl.plan_qpos_batch(pts[:1].repeat(4, 0), pts[1:2].repeat(4, 0), planning_time=10.0)
    assert all(r["status"] == "Success" for r in out) and stats["n_solved"] == 4


def test_ik_clik_parity_and_planner_ik():
    """batched computeIKCLIK (FP64 kernel) vs the oracle's restatement, then Planner.IK / plan."""
    from oracle import oracle as O

    pl = _planner()
    pm = pl.pinocchio_model
    li = pl.link_name_2_idx["panda_hand"]
    n = 512
    q_true = np.concatenate([random_q(n, 81), np.zeros((n, 2))], axis=1)
    poses = pm.compute_forward_kinematics_batch(q_true)[:, li, :].astype(np.float64)
    q0 = np.concatenate([random_q(n, 82), np.zeros((n, 2))], axis=1)
    mask = np.array([False] * 7 + [True] * 2)
    q, ok, err = pm.compute_IK_CLIK_batch(li, poses, q0, mask)
    desc = pm._device_scene().desc
    oq, ook, oerr = O.ik_clik_batch(desc, int(pm.link_user2model[li]), poses, q0, mask)
    print(f"CLIK success gpu={ok.mean():.3f} oracle={ook.mean():.3f}")
    # From a random start the damped iteration (damp 1e-12) passes near singular configurations and amplifies the last
    # bits of sincos / fma contraction, so the two double-precision runs may settle on different IK branches: compare
    # what the reference promises (success, and the pose is reached), and compare iterates where the problem is
    # well conditioned (start near the solution).
    assert ok.mean() > 0.9 and abs(ok.mean() - ook.mean()) < 0.02
    qn0 = q_true + np.concatenate([np.random.default_rng(83).normal(0, 0.1, (n, 7)), np.zeros((n, 2))], axis=1)
    qn, okn, _ = pm.compute_IK_CLIK_batch(li, poses, qn0, mask)
    oqn, ookn, _ = O.ik_clik_batch(desc, int(pm.link_user2model[li]), poses, qn0, mask)
    assert okn.all() and ookn.all() and np.abs(qn - oqn).max() < 1e-6
    assert np.linalg.norm(err[ok], axis=1).max() < 1e-5 and np.allclose(q[:, 7:], 0.0)
    reached = pm.compute_forward_kinematics_batch(q[ok])[:, li, :3]
    assert np.abs(reached - poses[ok, :3]).max() < 5e-5
    # single-problem reference signature
    q1, ok1, err1 = pm.compute_IK_CLIK(li, poses[0], q0[0], list(mask))
    assert ok1 == ok[0] and np.allclose(q1, q[0])
    with pytest.raises(RuntimeError):
        pm.compute_IK_CLIK(99, poses[0], q0[0])
    # Planner.IK: collision-free, within joint limits, reaches the pose
    start = np.array([0, 0.2, 0, -2.6, 0, 3.0, 0.8, 0, 0])
    goal = pm.compute_forward_kinematics_batch(np.array([[0.4, 0.3, -0.2, -2.0, 0.1, 2.4, 1.0, 0, 0]]))[0, li].astype(np.float64)
    status, q_goals = pl.IK(goal, start, mask, seed=5)
    assert status == "Success" and len(q_goals) >= 1
    for qg in q_goals:
        assert not pl.check_for_collision_batch(qg[None, :7])[0]
        got = pm.compute_forward_kinematics_batch(qg[None, :])[0, li]
        assert pl.distance_6D(goal[:3], goal[3:], got[:3].astype(np.float64), got[3:].astype(np.float64)) < 1e-3
    res = pl.plan(goal, start, mask, planning_time=10.0, seed=3)
    assert res["status"] == "Success" and np.allclose(res["position"][0], start[:7], atol=1e-6)
    # an unreachable pose fails with the reference's message
    far = goal.copy()
    far[:3] = [3.0, 0.0, 0.5]
    status, q_goals = pl.IK(far, start, mask, seed=5)
    assert status.startswith("IK Failed") and q_goals is None


def test_link_jacobian_parity_and_plan_screw():
    """batched getLinkJacobian (FP64 kernel) vs the oracle; Planner.plan_screw against a scalar
    restatement of mplib/planner.py:526-671 that uses the oracle's Jacobian and the oracle's collision check."""
    from oracle import oracle as O

    pl = _planner()
    pm = pl.pinocchio_model
    li = pl.link_name_2_idx["panda_hand"]
    desc = pm._device_scene().desc
    ml = int(pm.link_user2model[li])
    n = 1024
    q = np.concatenate([random_q(n, 91), np.random.default_rng(92).uniform(0, 0.04, (n, 2))], axis=1)
    for local in (False, True):
        J = pm.get_link_jacobian_batch(li, q, local)
        oJ, oT = O.link_jacobian_batch(desc, ml, q, local)
        assert J.shape == (n, 6, 9) and np.abs(J - oJ).max() < 1e-12
        assert np.all(J[:, :, 7:] == 0.0)  # the fingers do not move the hand
    _, T = pm._device_scene().link_jacobian(ml, q)
    assert np.abs(T - oT).max() < 1e-12
    # reference call sequence
    pm.compute_full_jacobian(q[0])
    assert np.allclose(pm.get_link_jacobian(li, local=False), O.link_jacobian_batch(desc, ml, q[:1], False)[0][0])
    assert np.allclose(pm.compute_single_link_local_jacobian(q[0], li), O.link_jacobian_batch(desc, ml, q[:1], True)[0][0])
    with pytest.raises(RuntimeError):
        pm.get_link_jacobian(99)

    # plan_screw: move the hand 15 cm along a straight line and rotate it 0.3 rad about z
    start = np.array([0, 0.2, 0, -2.6, 0, 3.0, 0.8, 0, 0])
    p0 = pm.compute_forward_kinematics_batch(start[None, :])[0, li].astype(np.float64)
    goal = p0.copy()
    goal[:3] += [0.05, 0.10, -0.10]
    cz, sz = np.cos(0.15), np.sin(0.15)
    w, x, y, z = p0[3:]
    goal[3:] = [cz * w - sz * z, cz * x - sz * y, cz * y + sz * x, cz * z + sz * w]  # Rz(0.3) * R0
    res = pl.plan_screw(goal, start, qpos_step=0.05)
    assert res["status"] == "Success"
    path = res["position"]
    assert path.shape[1] == 7 and np.allclose(path[0], start[:7])
    assert np.abs(np.linalg.norm(np.diff(path, axis=0), axis=1)[:-1] - 0.05).max() < 1e-9
    end = pm.compute_forward_kinematics_batch(np.concatenate([path[-1], start[7:]])[None, :])[0, li].astype(np.float64)
    assert pl.distance_6D(goal[:3], goal[3:], end[:3], end[3:]) < 2e-2
    assert not pl.check_for_collision_batch(path).any()

    # scalar restatement with the oracle underneath
    from mplib_b200.urdf import quat_to_mat

    cur, tgt = np.eye(4), np.eye(4)
    cur[:3, :] = O.link_jacobian_batch(desc, ml, start[None, :])[1][0]
    tgt[:3, 3] = goal[:3]
    tgt[:3, :3] = quat_to_mat(*goal[3:])
    tw, theta = pl._pose2exp_coordinate(tgt @ np.linalg.inv(cur))
    omega = tw * theta
    qq, ref = start.copy(), [start[:7].copy()]
    for _ in range(1000):
        Jr = O.link_jacobian_batch(desc, ml, qq[None, :])[0][0]
        dq = np.linalg.pinv(Jr) @ omega
        dq *= 0.05 / np.linalg.norm(dq)
        dtw = Jr @ dq
        flag = np.linalg.norm(dtw) > np.linalg.norm(omega)
        if flag:
            r = np.linalg.norm(omega) / np.linalg.norm(dtw)
            dq, dtw = dq * r, dtw * r
        qq += dq
        omega -= dtw
        ref.append(qq[:7].copy())
        if flag:
            break
    assert len(ref) == len(path) and np.abs(np.vstack(ref) - path).max() < 1e-9

    # a batch in lock step; with a floor below the robot a straight line down is refused at the first colliding waypoint
    from tests.workloads import box_surface_points

    pl.update_point_cloud(box_surface_points(np.random.Generator(np.random.PCG64(100)), 10000, (2.0, 2.0, 0.1), (0.0, 0.0, -0.1)), 1e-3)
    down = p0.copy()
    down[2] = -0.3
    out = pl.plan_screw_batch(np.stack([goal, down, goal]), np.stack([start, start, start]), qpos_step=0.05)
    assert out[0]["status"] == out[2]["status"] == "Success" and np.allclose(out[0]["position"], path)
    assert out[1]["status"] == "screw plan failed"
