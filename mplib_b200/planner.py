"""Host mirror of ``mplib.Planner`` (mplib/planner.py) for the calls that sit on the accelerated path: collision
queries, SRDF auto-generation, IK, joint-space planning and screw motions.  Every numeric query goes through the C-ABI
to the GPU.

Out of scope here: TOPPRA time
parametrisation (the `time` / `velocity` / `acceleration` keys of the reference's result dict) -- `toppra` is absent
from this image; `plan_qpos` returns the waypoint matrix under "position" exactly as the reference hands it to TOPPRA
(mplib/planner.py:493-510)."""
from __future__ import annotations

import os
import pathlib
import xml.etree.ElementTree as ET
from xml.dom import minidom

import numpy as np

from .pymp import ArticulatedModel, OMPLPlanner, PlanningWorld
from .urdf import quat_to_mat as _quat_to_mat


class Planner:
    def __init__(self, urdf, user_link_names, user_joint_names, move_group, joint_vel_limits=None, joint_acc_limits=None, srdf="",
                 package_keyword_replacement="", device=0, srdf_sample_time=1 << 20):
        """mplib/planner.py:19-97 (same argument names and order)."""
        self.urdf = urdf
        if srdf == "" and os.path.exists(urdf.replace(".urdf", ".srdf")):
            srdf = urdf.replace(".urdf", ".srdf")
            print("No SRDF file provided. Try to load %s." % srdf)
        self.srdf = srdf
        self.user_link_names = list(user_link_names)
        self.user_joint_names = list(user_joint_names)
        self.joint_name_2_idx = {j: i for i, j in enumerate(self.user_joint_names)}
        self.link_name_2_idx = {l: i for i, l in enumerate(self.user_link_names)}
        urdf = self.replace_package_keyword(package_keyword_replacement)
        self.robot = ArticulatedModel(urdf, srdf, [0, 0, -9.81], self.user_joint_names, self.user_link_names, verbose=False, convex=True)
        self.pinocchio_model = self.robot.get_pinocchio_model()
        self.planning_world = PlanningWorld([self.robot], ["robot"], [], [], device=device)
        self.acm = self.planning_world.get_allowed_collision_matrix()
        if srdf == "":
            self.generate_collision_pair(sample_time=srdf_sample_time)
            self.robot.update_SRDF(self.srdf)
        assert move_group in self.user_link_names
        self.move_group = move_group
        self.robot.set_move_group(self.move_group)
        self.move_group_joint_indices = self.robot.get_move_group_joint_indices()
        self.joint_types = self.pinocchio_model.get_joint_types()
        self.joint_limits = np.concatenate(self.pinocchio_model.get_joint_limits())
        self.planner = OMPLPlanner(world=self.planning_world)
        self.joint_vel_limits = joint_vel_limits
        self.joint_acc_limits = joint_acc_limits
        self.move_group_link_id = self.link_name_2_idx[self.move_group]

    def replace_package_keyword(self, package_keyword_replacement):
        """Behaviour of mplib/planner.py:99-116: a URDF that refers to ``package://`` resources is written once, next to the
        original, with that prefix replaced; the path of the file to load is returned."""
        text = pathlib.Path(self.urdf).read_text()
        if "package://" not in text:
            return self.urdf
        stem, ext = os.path.splitext(self.urdf)
        patched = pathlib.Path(stem + "_package_keyword_replaced" + (ext or ".urdf"))
        if not patched.exists():
            patched.write_text(text.replace("package://", package_keyword_replacement))
        return str(patched)

    # ------------------------------------------------------------------ SRDF auto-generation
    def always_colliding_pairs(self, sample_time=1 << 20, seed=None, batch=1 << 18):
        """The statistic of `generate_collision_pair` (mplib/planner.py:118-163): how often every admitted link pair
        collides over `sample_time` random configurations (uniform in the joint limits, as
        `get_random_configuration`), as ONE batched collide_full per chunk: the pair masks come back as bit
        columns and are summed per pair.  -> (count[n_link, n_link] int64, samples)"""
        w = self.planning_world
        rng = np.random.Generator(np.random.PCG64(seed)) if seed is not None else np.random.default_rng()
        old_qpos = self.robot.get_qpos().copy()
        old_ee = self.robot.get_move_group_end_effectors() if getattr(self.robot, "move_group_end_effectors", None) else None
        # the reference samples ALL joints (set_qpos(qpos, full=True)): every joint becomes a batch state for the statistic
        self.robot.set_move_group(self.pinocchio_model.get_leaf_links())
        idx = self.robot.get_move_group_joint_indices()
        lim = np.concatenate(self.pinocchio_model.get_joint_limits())[idx]
        lo, hi = lim[:, 0], lim[:, 1]
        n_link = len(self.user_link_names)
        cnt = np.zeros((n_link, n_link), dtype=np.int64)
        done = 0
        while done < sample_time:
            m = min(batch, sample_time - done)
            cnt += self.pair_collision_counts(lo + (hi - lo) * rng.random((m, len(idx))))
            done += m
        if old_ee:
            self.robot.set_move_group(old_ee)
        self.robot.set_qpos(old_qpos, True)
        return cnt, sample_time

    def pair_collision_counts(self, states):
        """count[u, v] = number of `states` in which link u collides with link v (one batched collide_full)."""
        w = self.planning_world
        n_link = len(self.user_link_names)
        cnt = np.zeros((n_link, n_link), dtype=np.int64)
        _, ph = w.collide_full_batch(states, want_pairs=True)
        meta = w.get_pair_metadata()
        for p, c in enumerate(ph.sum(axis=0)):
            if c and meta[p][4] in self.link_name_2_idx and meta[p][5] in self.link_name_2_idx:
                cnt[self.link_name_2_idx[meta[p][4]], self.link_name_2_idx[meta[p][5]]] += int(c)
        return cnt

    def generate_collision_pair(self, sample_time=1000000, echo_freq=100000, seed=None):
        """mplib/planner.py:118-163: link pairs that collide in EVERY sample are written to <urdf>.srdf as
        disable_collisions entries."""
        print("Since no SRDF file is provided, we will first detect link pairs that will always collide.")
        cnt, n = self.always_colliding_pairs(sample_time, seed)
        root = ET.Element("robot")
        root.set("name", self.urdf.split("/")[-1].split(".")[0])
        self.srdf = self.urdf.replace(".urdf", ".srdf")
        for i in range(cnt.shape[0]):
            for j in range(cnt.shape[1]):
                if cnt[i][j] == n:
                    link1, link2 = self.user_link_names[i], self.user_link_names[j]
                    print("Ignore collision pair: (%s, %s), reason:  always collide" % (link1, link2))
                    c = ET.SubElement(root, "disable_collisions")
                    c.set("link1", link1)
                    c.set("link2", link2)
                    c.set("reason", "Default")
        with open(self.srdf, "w") as f:
            f.write(minidom.parseString(ET.tostring(root)).toprettyxml(indent="    "))
        print("Saving the SRDF file to %s" % self.srdf)

    # ------------------------------------------------------------------ collision queries
    def check_joint_limit(self, qpos):
        """mplib/planner.py:170-192 (qpos is clipped in place for revolute joints)."""
        for i, (q, joint_type, qlimit) in enumerate(zip(qpos, self.joint_types, self.joint_limits)):
            if joint_type.startswith("JointModelR"):
                if np.abs(q - qlimit[0]) < 1e-3:
                    continue
                q -= 2 * np.pi * np.floor((q - qlimit[0]) / (2 * np.pi))
                qpos[i] = q
                if q > qlimit[1] + 1e-3:
                    return False
            elif q < qlimit[0] - 1e-3 or q > qlimit[1] + 1e-3:
                return False
        return True

    def check_for_collision(self, collision_function, state=None):
        """Behaviour of mplib/planner.py:194-224: run ``collision_function`` with the (single) planned articulation
        placed at ``state`` -- a full qpos or just the move-group joints, default: where it is -- and put it back."""
        arts = self.planning_world.get_planned_articulations()
        if len(arts) != 1:
            raise AssertionError("Only support 1 planned articulation now")
        art = arts[0]
        saved = art.get_qpos()
        q = saved.copy()
        if state is not None:
            given = np.asarray(state, dtype=np.float64).reshape(-1)
            if given.shape[0] == len(self.move_group_joint_indices):
                q[: given.shape[0]] = given  # move-group joints come first in the user joint order
            else:
                q = given
        art.set_qpos(q, True)
        try:
            return collision_function()
        finally:
            art.set_qpos(saved, True)

    def check_for_self_collision(self, qpos=None):
        """mplib/planner.py:226-240"""
        return self.check_for_collision(self.planning_world.self_collide, qpos)

    def check_for_env_collision(self, qpos=None):
        """mplib/planner.py:242-256"""
        return self.check_for_collision(self.planning_world.collide_with_others, qpos)

    def check_for_collision_batch(self, states):
        """Batched form: states [N, len(move group joints)] -> colliding bool[N] (self + environment)."""
        return self.planning_world.collide_batch(np.asarray(states, dtype=np.float64))

    # ------------------------------------------------------------------ inverse kinematics
    def distance_6D(self, p1, q1, p2, q2):
        """mplib/planner.py:165-168"""
        return np.linalg.norm(p1 - p2) + min(np.linalg.norm(q1 - q2), np.linalg.norm(q1 + q2))

    def IK(self, goal_pose, start_qpos, mask=(), *, n_init_qpos=20, threshold=0.001, return_closest=False, verbose=False, seed=None):
        """mplib/planner.py:256-356.  The reference runs its `n_init_qpos` CLIK restarts, the joint-limit check, the
        collision check and the pose check one after the other; here all restarts are ONE batched CLIK call, one batched
        collide and one batched FK, then the same acceptance logic in the same order."""
        status, goals = self.IK_batch(np.asarray(goal_pose, dtype=np.float64)[None, :], start_qpos, mask, n_init_qpos=n_init_qpos,
                                      threshold=threshold, return_closest=return_closest, seed=seed)
        return status[0], goals[0]

    def IK_batch(self, goal_poses, start_qpos, mask=(), *, n_init_qpos=20, threshold=0.001, return_closest=False, seed=None):
        """`IK` for G goal poses at once ([G, 7]); one start_qpos for all.  -> (list of status, list of q_goals)"""
        goal_poses = np.asarray(goal_poses, dtype=np.float64).reshape(-1, 7)
        start_qpos = np.asarray(start_qpos, dtype=np.float64)
        G, ndof = goal_poses.shape[0], start_qpos.shape[0]
        mask = np.zeros(ndof, dtype=bool) if len(mask) == 0 else np.asarray(mask, dtype=bool)
        move_link_idx = self.link_name_2_idx[self.move_group]
        move_joint_idx = self.move_group_joint_indices
        rng = np.random.Generator(np.random.PCG64(seed)) if seed is not None else np.random.default_rng()
        lim = np.concatenate(self.pinocchio_model.get_joint_limits())
        # restart 0 starts from start_qpos, the others from random configurations with the masked joints kept
        q0 = lim[:, 0] + (lim[:, 1] - lim[:, 0]) * rng.random((G, n_init_qpos, ndof))
        q0[:, 0, :] = start_qpos
        q0[:, :, mask] = start_qpos[mask]
        poses_rep = np.repeat(goal_poses, n_init_qpos, axis=0)
        q, ok, _ = self.pinocchio_model.compute_IK_CLIK_batch(move_link_idx, poses_rep, q0.reshape(-1, ndof), mask)
        for k in range(q.shape[0]):  # check_joint_limit() clips qpos towards the lower limits
            ok[k] = ok[k] and self.check_joint_limit(q[k])
        self.robot.set_qpos(start_qpos, True)
        if ok.any():
            cand = np.nonzero(ok)[0]
            ok[cand[self.planning_world.collide_batch(q[cand][:, move_joint_idx])]] = False
        dist = np.full(q.shape[0], np.inf)
        if ok.any():
            cand = np.nonzero(ok)[0]
            new_pose = self.pinocchio_model.compute_forward_kinematics_batch(q[cand])[:, move_link_idx, :].astype(np.float64)
            for c, npz in zip(cand, new_pose):
                g = poses_rep[c]
                dist[c] = self.distance_6D(g[:3], g[3:], npz[:3], npz[3:])
        status, out = [], []
        for gi in range(G):
            sl = slice(gi * n_init_qpos, (gi + 1) * n_init_qpos)
            q_goals, min_dis = [], 1e9
            for k in np.nonzero(ok[sl])[0] + gi * n_init_qpos:
                min_dis = min(min_dis, dist[k])
                if dist[k] < threshold and not any(np.linalg.norm(qg[move_joint_idx] - q[k][move_joint_idx]) < 0.1 for qg in q_goals):
                    q_goals.append(q[k].copy())
            if q_goals:
                if return_closest:
                    q_goals = self._closest_goal(np.asarray(q_goals), start_qpos)
                status.append("Success")
                out.append(q_goals)
            elif min_dis != 1e9:
                status.append(f"IK Failed! Distance {min_dis} is greater than {threshold=}.")
                out.append(None)
            else:
                status.append("IK Failed! Cannot find valid solution.")
                out.append(None)
        return status, out

    def _closest_goal(self, q_goals, start_qpos):
        """mplib/planner.py:337-355: the solution closest to start_qpos, considering +2*pi images of wide revolute joints"""
        equiv = np.array([t.startswith("JointModelR") for t in self.joint_types]) & (self.joint_limits[:, 1] - self.joint_limits[:, 0] > 2 * np.pi)
        start = np.asarray(start_qpos)[None]
        q1 = q_goals[:, equiv]
        q2 = q1 + 2 * np.pi
        closer = (q2 < self.joint_limits[:, 1][None, equiv]) & (np.abs(q1 - start[:, equiv]) > np.abs(q2 - start[:, equiv]))
        q_goals[:, equiv] = np.where(closer, q2, q1)
        return q_goals[np.linalg.norm(q_goals - start, axis=1).argmin()]

    def plan(self, goal_pose, current_qpos, mask=(), *, planner_name="RRTConnect", rrt_range=0.1, planning_time=1.0,
             fix_joint_limits=True, verbose=False, **kw):
        """mplib/planner.py:415-524 without the TOPPRA step: IK for the goal pose, then RRTConnect to the IK solutions."""
        current_qpos = np.array(current_qpos, dtype=np.float64, copy=True)
        if fix_joint_limits:
            current_qpos = np.clip(current_qpos, self.joint_limits[:, 0], self.joint_limits[:, 1])
        ik_status, goal_qpos = self.IK(goal_pose, current_qpos, mask)
        if ik_status != "Success":
            return {"status": ik_status}
        return self.plan_qpos(goal_qpos, current_qpos, planner_name=planner_name, rrt_range=rrt_range, planning_time=planning_time,
                              fix_joint_limits=False, verbose=verbose, **kw)

    # ------------------------------------------------------------------ scene updates (mplib/planner.py:358-413)
    def update_point_cloud(self, pc, radius=1e-3):
        self.planning_world.add_point_cloud("scene_pcd", np.asarray(pc, dtype=np.float64), radius)

    def remove_point_cloud(self):
        return self.planning_world.remove_normal_object("scene_pcd")

    def update_attached_sphere(self, radius, pose, art_name="robot", link_id=-1):
        self.planning_world.attach_sphere(radius, art_name, self.move_group_link_id if link_id == -1 else link_id, pose)

    def update_attached_box(self, size, pose, art_name="robot", link_id=-1):
        self.planning_world.attach_box(size, art_name, self.move_group_link_id if link_id == -1 else link_id, pose)

    def update_attached_mesh(self, mesh_path, pose, art_name="robot", link_id=-1):
        self.planning_world.attach_mesh(mesh_path, art_name, self.move_group_link_id if link_id == -1 else link_id, pose)

    def detach_object(self, name="attached_geom", also_remove=False):
        return self.planning_world.detach_object(name, also_remove)

    # ------------------------------------------------------------------ joint-space planning
    def plan_qpos(self, goal_qposes, current_qpos, *, planner_name="RRTConnect", rrt_range=0.1, planning_time=1.0,
                  fix_joint_limits=True, verbose=False, **kw):
        """RRTConnect from `current_qpos` to any of `goal_qposes` (the RRT half of `Planner.plan`,
        mplib/planner.py:455-524, after its IK step).  -> {"status", "position"[n, ndof]}"""
        current_qpos = np.array(current_qpos, dtype=np.float64, copy=True)
        n = current_qpos.shape[0]
        if fix_joint_limits:
            for i in range(n):
                current_qpos[i] = np.clip(current_qpos[i], self.joint_limits[i][0], self.joint_limits[i][1])
        self.robot.set_qpos(current_qpos, True)
        if self.planning_world.collide():
            if verbose:
                print("Invalid start state!")
        idx = self.move_group_joint_indices
        goals = [np.asarray(g, dtype=np.float64)[idx] if len(g) == n and n != len(idx) else np.asarray(g, dtype=np.float64)[: len(idx)]
                 for g in goal_qposes]
        status, path = self.planner.plan(current_qpos[idx], goals, planner_name=planner_name, time=planning_time, range=rrt_range,
                                         verbose=verbose, **kw)
        if status == "Exact solution":
            return {"status": "Success", "position": path}
        return {"status": "%s failed. %s" % (planner_name, status)}

    # ------------------------------------------------------------------ screw motion
    @staticmethod
    def _pose2exp_coordinate(T):
        """mplib/planner.py:574-604: (twist direction [v, w], angle) of a 4x4 rigid transform."""
        R = T[:3, :3]
        tr = R.trace()
        if np.isclose(tr, 3):
            omega, theta = np.zeros(3), 1
        elif np.isclose(tr, -1):
            return np.zeros(3), -1e6
        else:
            theta = np.arccos((tr - 1) / 2)
            omega = 1 / 2 / np.sin(theta) * np.array([R[2, 1] - R[1, 2], R[0, 2] - R[2, 0], R[1, 0] - R[0, 1]])
        ss = np.array([[0, -omega[2], omega[1]], [omega[2], 0, -omega[0]], [-omega[1], omega[0], 0]])
        inv_left_jacobian = np.eye(3) / theta - 0.5 * ss + (1.0 / theta - 0.5 / np.tan(theta / 2)) * ss @ ss
        return np.concatenate([inv_left_jacobian @ T[:3, 3], omega]), theta

    def plan_screw_batch(self, goal_poses, current_qposes, *, qpos_step=0.1, max_steps=100000):
        """Q straight-line screw motions in lock step (mplib/planner.py:526-671 per query).  Every step evaluates the
        WORLD Jacobian of the move-group link for all live queries in one device call (mpb_link_jacobian_batch), takes
        the pseudo-inverse step of norm `qpos_step` on the host, and checks all new states with one mpb_collide_batch.
        Joints outside the move group keep the articulation's current values in the collision check.
        -> list of {"status", "position"[n, len(move group joints)]}; TOPPRA is not run (the waypoints are returned)."""
        goal_poses = np.asarray(goal_poses, dtype=np.float64).reshape(-1, 7)
        q = np.array(current_qposes, dtype=np.float64, copy=True)
        Q, nv = q.shape
        assert goal_poses.shape[0] == Q
        idx = np.asarray(self.move_group_joint_indices)
        pm = self.pinocchio_model
        dev, link = pm._device_scene(), int(pm.link_user2model[self.link_name_2_idx[self.move_group]])
        _, T0 = dev.link_jacobian(link, q)
        omega = np.zeros((Q, 6))
        live = np.ones(Q, dtype=bool)
        results = [None] * Q
        for i in range(Q):
            cur, tgt = np.eye(4), np.eye(4)
            cur[:3, :] = T0[i]
            tgt[:3, 3] = goal_poses[i, :3]
            tgt[:3, :3] = np.asarray(_quat_to_mat(*goal_poses[i, 3:])).reshape(3, 3)
            w, theta = self._pose2exp_coordinate(tgt @ np.linalg.inv(cur))
            if theta < -1e4:
                results[i], live[i] = {"status": "screw plan failed."}, False
            omega[i] = w * theta
        paths = [[q[i, idx].copy()] for i in range(Q)]
        lo, hi = self.joint_limits[:, 0] - 1e-3, self.joint_limits[:, 1] + 1e-3
        for _ in range(max_steps):
            a = np.nonzero(live)[0]
            if a.size == 0:
                break
            J, _ = dev.link_jacobian(link, q[a])
            dq = np.einsum("nij,nj->ni", np.linalg.pinv(J), omega[a])
            with np.errstate(divide="ignore", invalid="ignore"):
                dq *= (qpos_step / np.linalg.norm(dq, axis=1))[:, None]
            dtw = np.einsum("nij,nj->ni", J, dq)
            n_tw, n_om = np.linalg.norm(dtw, axis=1), np.linalg.norm(omega[a], axis=1)
            flag = n_tw > n_om
            with np.errstate(divide="ignore", invalid="ignore"):
                ratio = np.where(flag, n_om / n_tw, 1.0)
            dq, dtw = dq * ratio[:, None], dtw * ratio[:, None]
            q[a] += dq
            omega[a] -= dtw
            within = np.all((q[a] >= lo) & (q[a] <= hi), axis=1)
            collide = self.planning_world.collide_batch(q[a][:, idx])
            bad = ~(np.linalg.norm(dtw, axis=1) >= 1e-4) | collide | ~within
            for k, i in enumerate(a):
                if bad[k]:
                    results[i], live[i] = {"status": "screw plan failed"}, False
                    continue
                paths[i].append(q[i, idx].copy())
                if flag[k]:
                    results[i], live[i] = {"status": "Success", "position": np.vstack(paths[i])}, False
        for i in np.nonzero(live)[0]:
            results[i] = {"status": "screw plan failed"}
        return results

    def plan_screw(self, goal_pose, current_qpos, *, time_step=0.1, qpos_step=0.1, verbose=False):
        """mplib/planner.py:526-671 without the TOPPRA step (toppra is not in this image: the waypoints are returned
        as "position", the hand-off format of mplib/planner.py:358-375)."""
        current_qpos = np.array(current_qpos, dtype=np.float64, copy=True)
        self.robot.set_qpos(current_qpos, True)
        return self.plan_screw_batch(np.asarray(goal_pose, dtype=np.float64)[None, :], current_qpos[None, :], qpos_step=qpos_step)[0]

    def plan_qpos_batch(self, start_qposes, goal_qposes, *, rrt_range=0.1, planning_time=1.0, **kw):
        """Q independent joint-space queries at once (see OMPLPlanner.plan_batch).  -> list of result dicts"""
        status, paths, stats = self.planner.plan_batch(np.asarray(start_qposes, dtype=np.float64), np.asarray(goal_qposes, dtype=np.float64),
                                                       time=planning_time, range=rrt_range, **kw)
        out = [{"status": "Success", "position": p} if s == "Exact solution" else {"status": "RRTConnect failed. %s" % s}
               for s, p in zip(status, paths)]
        return out, stats
