"""Shared builders for the tests: the Panda world of the reference's own tests
(tests/test_basic.py:5-36) and seeded random configurations."""
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
URDF = os.path.join(ROOT, "data/panda/panda.urdf")
SRDF = os.path.join(ROOT, "data/panda/panda.srdf")

PANDA_LINKS = ["panda_link0", "panda_link1", "panda_link2", "panda_link3", "panda_link4", "panda_link5", "panda_link6",
               "panda_link7", "panda_link8", "panda_hand", "panda_leftfinger", "panda_rightfinger"]
PANDA_JOINTS = ["panda_joint1", "panda_joint2", "panda_joint3", "panda_joint4", "panda_joint5", "panda_joint6",
                "panda_joint7", "panda_finger_joint1", "panda_finger_joint2"]
PANDA_LO = np.array([-2.8973, -1.7628, -2.8973, -3.0718, -2.8973, -0.0175, -2.8973])
PANDA_HI = np.array([2.8973, 1.7628, 2.8973, -0.0698, 2.8973, 3.7525, 2.8973])


def make_panda(convex=True, urdf=URDF, srdf=SRDF):
    from mplib_b200.pymp import ArticulatedModel

    art = ArticulatedModel(urdf, srdf, (0, 0, -9.81), PANDA_JOINTS, PANDA_LINKS, verbose=False, convex=convex)
    art.set_move_group("panda_hand")
    return art


def make_panda_world(convex=True):
    from mplib_b200.pymp import PlanningWorld

    return PlanningWorld([make_panda(convex)], ["robot"])


def random_q(n, seed, lo=PANDA_LO, hi=PANDA_HI):
    rng = np.random.Generator(np.random.PCG64(seed))
    return lo + (hi - lo) * rng.random((n, lo.shape[0]))


def compile_desc(world):
    """World -> SceneDesc without touching the GPU (for oracle-only tests)."""
    b, cand = world.compile()
    for p in world._admitted(cand):
        b.add_pair(p[0], p[1], p[3])
    return b.finish()


def has_gpu():
    try:
        from mplib_b200 import _lib

        return _lib.load().mpb_device_count() > 0
    except Exception:
        return False
