"""``pymp.ompl.OMPLPlanner`` mirror (``src/ompl_planner.{h,cpp}``, ``python/pybind_ompl.hpp:25-30``).

Host logic only: the state-space construction, the start fix-up and the goal-image enumeration of
``OMPLPlannerTpl::plan`` run here; validity checks, motion validation and RRTConnect itself run on the GPU through
``mpb_collide_batch`` / ``mpb_check_motion_batch`` / ``mpb_rrtconnect_batch``.  ``plan`` keeps the reference signature
and return value ``(status, path[n, dim])``; ``plan_batch`` / ``check_motion_batch`` are the additive batch forms.

Deviations (documented in DESIGN.md): only ``RRTConnect`` is ported (the path BASELINE.json names); the RNG differs
from OMPL's, so paths are valid but not identical; a query that runs out of budget returns ``("Timeout", empty)``
instead of OMPL's approximate start-tree path.
"""
from __future__ import annotations

import builtins

import numpy as np

PI = 3.14159265359  # src/ompl_planner.cpp: #define PI 3.14159265359

_STATUS = {0: "Timeout", 1: "Exact solution", 2: "Timeout", 3: "Invalid goal", 4: "Exact solution"}
_global_rng = np.random.RandomState(0)


def set_global_seed(seed):
    """``set_global_seed`` (``python/pybind_planning_world.hpp:32``, ``src/random_utils.h:12``)."""
    global _global_rng
    _global_rng = np.random.RandomState(int(seed) & 0x7FFFFFFF)
    np.random.seed(int(seed) & 0x7FFFFFFF)


class OMPLPlanner:
    def __init__(self, world):
        self.world = world
        self.build_state_space()

    def build_state_space(self):
        """``src/ompl_planner.cpp:245-296``: planned articulations in name order, move-group joints only."""
        lo, hi, so2, rev = [], [], [], []
        for art in self.world.get_planned_articulations():
            pm = art.get_pinocchio_model()
            types = pm.get_joint_types()
            limits = pm.get_joint_limits()
            d = 0
            for jid in art.get_move_group_joint_indices():
                t = types[jid][len("JointModel"):]
                if t[0] == "P" or (t[0] == "R" and t[1:2] != "U"):
                    b = limits[jid]
                    for row in b:
                        lo.append(float(row[0])); hi.append(float(row[1])); so2.append(0); d += 1
                    rev.append(t[0] == "R")
                elif t[:2] == "RU":
                    lo.append(-PI); hi.append(PI); so2.append(1); rev.append(False); d += 1
            if d != art.get_qpos_dim():
                raise RuntimeError(f"Dim of bound is different from dim of qpos {d} {art.get_qpos_dim()}")
        self.lower_joint_limits = np.asarray(lo, dtype=np.float64)
        self.upper_joint_limits = np.asarray(hi, dtype=np.float64)
        self.is_revolute = rev
        self.is_so2 = np.asarray(so2, dtype=np.uint8)
        self.dim = len(lo)
        from .._lib import Space

        self.space = Space(self.lower_joint_limits, self.upper_joint_limits, self.is_so2)

    # -- pieces of OMPLPlannerTpl::plan ----------------------------------------------------------------
    def _is_valid(self, states):
        return ~self.world.collide_batch(np.atleast_2d(np.asarray(states, dtype=np.float64)))

    def random_sample_nearby(self, start_state):
        """``src/ompl_planner.cpp:72-95``: the first valid of up to 1001 growing perturbations (all candidates are
        drawn up front and validated in one batch)."""
        start_state = np.asarray(start_state, dtype=np.float64)
        ratio = (np.arange(1001) + 1) / 1000.0
        r = _global_rng.random_sample((1001, self.dim)) * 2 - 1
        cand = start_state[None, :] + (self.upper_joint_limits - self.lower_joint_limits)[None, :] * ratio[:, None] * r
        cand = np.clip(cand, self.lower_joint_limits, self.upper_joint_limits)
        ok = np.nonzero(self._is_valid(cand))[0]
        if ok.size:
            print(f"successfully sampled a new state with a perturbation of {ratio[ok[0]] * 100}% joint limits.")
            return cand[ok[0]]
        return start_state

    def goal_images(self, goal_states):
        """``src/ompl_planner.cpp:118-158``: every +-2*pi image of the revolute coordinates inside the limits."""
        out = []
        tot = 3 ** self.dim
        for g in goal_states:
            g = np.asarray(g, dtype=np.float64).reshape(-1)
            # the 3^dim enumeration collapses to the product of the per-joint admissible shifts
            per_joint = []
            for j in range(self.dim):
                opts = [(0, g[j])]
                if self.is_revolute[j]:
                    if g[j] - 2 * PI > self.lower_joint_limits[j]:
                        opts.append((1, g[j] - 2 * PI))
                    if g[j] + 2 * PI < self.upper_joint_limits[j]:
                        opts.append((2, g[j] + 2 * PI))
                per_joint.append(opts)
            n_img = int(np.prod([len(o) for o in per_joint]))
            assert n_img <= tot
            # enumeration order of the reference: i = sum dir_j 3^j ascending
            imgs = [((), [])]
            for j in range(self.dim):
                imgs = [(code + (d,), vals + [v]) for code, vals in imgs for d, v in per_joint[j]]
            imgs.sort(key=lambda cv: sum(d * 3 ** j for j, d in enumerate(cv[0])))
            out.extend(np.asarray(v) for _, v in imgs)
        return out

    def plan(self, start_state, goal_states, planner_name="RRTConnect", time=1.0, range=0.0, goal_bias=0.05,
             pathlen_obj_weight=10.0, pathlen_obj_only=False, verbose=False, **kw):
        start_state = np.asarray(start_state, dtype=np.float64).reshape(-1)
        goal_states = [np.asarray(g, dtype=np.float64).reshape(-1) for g in goal_states]
        if start_state.shape[0] != goal_states[0].shape[0]:
            raise RuntimeError("Length of start state and goal state should be equal")
        if start_state.shape[0] != self.dim:
            raise RuntimeError("Length of start state and problem dimension should be equal")
        status, paths, _ = self.plan_batch(start_state[None, :], [goal_states], planner_name, time, range, verbose=verbose, **kw)
        return status[0], paths[0]

    def plan_batch(self, start_states, goal_states, planner_name="RRTConnect", time=1.0, range=0.0, verbose=False,
                   max_rounds=8192, tree_capacity=4096, max_path=2048, seed=None, check_every=16, connect_steps=8, n_slots=0,
                   max_rounds_per_query=0):
        """Q independent queries.  start_states [Q, dim]; goal_states: per query a list of goal states (or an array
        [Q, dim] for one goal each).  -> (list of status strings, list of float64 paths [n_i, dim], stats)
        ``time`` bounds the whole call; ``max_rounds_per_query`` is the per-query budget (the role ``time`` plays for the
        reference's one query per call); ``n_slots`` < Q runs the queries through that many concurrent searches, a
        finished query's slot taking the next pending one."""
        if planner_name != "RRTConnect":
            if planner_name in ("RRT", "PRMstar", "LazyPRMstar", "RRTstar", "RRTsharp", "RRTXstatic", "InformedRRTstar"):
                raise NotImplementedError(f"{planner_name} is outside the accelerated path (only RRTConnect is ported)")
            raise RuntimeError("Planner Not implemented")
        starts = np.array(np.asarray(start_states, dtype=np.float64), copy=True)
        Q = starts.shape[0]
        if isinstance(goal_states, np.ndarray) and goal_states.ndim == 2:
            goal_states = [[g] for g in goal_states]
        # start fix-up (src/ompl_planner.cpp:110-115)
        invalid_start = ~self._is_valid(starts)
        orig = starts.copy()
        for i in np.nonzero(invalid_start)[0]:
            print("invalid start state!! (collision)")
            starts[i] = self.random_sample_nearby(starts[i])
        # goal images; invalid goal states are skipped as pis_.nextGoal() does
        # fast path: a goal none of whose revolute coordinates has a +-2*pi image inside the limits is its own only image
        imgs = []
        for gs in goal_states:
            if len(gs) == 1:
                g = np.asarray(gs[0], dtype=np.float64).reshape(-1)
                if not (self.is_revolute & ((g - 2 * PI > self.lower_joint_limits) | (g + 2 * PI < self.upper_joint_limits))).any():
                    imgs.append([g])
                    continue
            imgs.append(self.goal_images(gs))
        flat = np.asarray([g for im in imgs for g in im], dtype=np.float64).reshape(-1, self.dim)
        ok = self._is_valid(flat) if len(flat) else np.zeros(0, bool)
        k = 0
        kept = []
        for im in imgs:
            kept.append([g for g, v in zip(im, ok[k:k + len(im)]) if v])
            k += len(im)
        G = max(1, max(len(x) for x in kept))
        goals = np.zeros((Q, G, self.dim), dtype=np.float32)
        ng = np.zeros(Q, dtype=np.int32)
        for i, x in enumerate(kept):
            ng[i] = len(x)
            for j, g in enumerate(x):
                goals[i, j] = g
        if verbose:
            print("number of goal state:", int(ng.sum()))
        dev = self.world._sync()["dev"]
        seed = int(_global_rng.randint(0, 2 ** 31 - 1)) if seed is None else int(seed)
        st, paths, stats = dev.rrtconnect(self.space, starts, goals, ng, range_=range, max_rounds=max_rounds, time_limit=time,
                                          tree_capacity=tree_capacity, max_path=max_path, seed=seed, check_every=check_every,
                                          connect_steps=connect_steps, n_slots=n_slots, max_rounds_per_query=max_rounds_per_query)
        status, out = [], []
        for i in builtins.range(Q):  # `range` is the reference's argument name
            s = _STATUS.get(int(st[i]), "Unknown")
            p = paths[i]
            if int(st[i]) == 1 and invalid_start[i]:
                p = np.concatenate([orig[i][None, :], p], axis=0)  # src/ompl_planner.cpp:226-230
            if int(st[i]) != 1:
                p = np.zeros((0, self.dim))
                if int(st[i]) == 4:
                    s = "Timeout"
            status.append(s)
            out.append(p)
        return status, out, stats

    def check_motion_batch(self, a, b):
        """OMPL's DiscreteMotionValidator for E edges at once: [E, dim] x [E, dim] -> valid bool[E]."""
        return self.world._sync()["dev"].check_motion(self.space, a, b)
