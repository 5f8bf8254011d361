"""Synthetic workloads of SURVEY 8(d) / BASELINE.json configs (seeds and shapes fixed there) and the Panda world of the
reference's own tests (tests/test_basic.py:5-36).  Used by bench.py and by tests/ (which re-export it)."""
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))  # the repository root (data/ lives there)
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




def _flop_model(desc, n_attached=0, n_robots=1):
    """F_fk = sum_joints(2+80) + 63*(#links + #objects + #attached); V_mean over admitted pairs."""
    f_fk = desc.n_joints * 82 + 63 * (desc.n_links + int((desc.obj_link >= 0).sum()))
    nv = np.diff(desc.hull_vert_off)
    vs = []
    for a, b in zip(desc.pair_a, desc.pair_b):
        for o in (a, b):
            vs.append(nv[desc.obj_geom[o]] if desc.obj_type[o] == 0 else 1)
    return {"F_fk": float(f_fk), "V_mean": float(np.mean(vs)) if vs else 0.0}


def c2(batch=0):
    """configs[1]: Panda batched FK + self-collision on convex hulls, 1M random configurations."""
    n = batch or (1 << 20)
    w = make_panda_world(convex=True)
    desc = compile_desc(w)
    return {
        "world": w, "desc": desc, "batch": n,
        "make_q": lambda k, m: random_q(m, 1 + 7919 * k).astype(np.float32),
        "config": {"workload": "configs[1]: Panda FK + self-collision, convex hulls, 19 ACM/SRDF-admitted pairs, "
                               f"{n} random configurations per GPU per step (uniform in URDF limits, seed 1)",
                   "states_per_gpu": n, "state_dim": desc.state_dim, "pairs": desc.n_pairs,
                   "outputs": "verdict bit + 19-bit pair mask per configuration",
                   "l2_policy": "rotating pool of input batches > 126 MB (no step re-reads cached q)"},
        "flop_model": _flop_model(desc),
    }




def _rand_rot(rng):
    """Uniform rotation (unit quaternion w,x,y,z) from 3 uniforms (Shoemake)."""
    u1, u2, u3 = rng.random(3)
    return np.array([np.sqrt(u1) * np.cos(2 * np.pi * u3), np.sqrt(1 - u1) * np.sin(2 * np.pi * u2),
                     np.sqrt(1 - u1) * np.cos(2 * np.pi * u2), np.sqrt(u1) * np.sin(2 * np.pi * u3)])


def _shell_point(rng):
    while True:
        p = rng.uniform(-1.0, 1.0, 3)
        r = np.linalg.norm(p)
        if 0.30 <= r <= 1.0 and p[2] >= 0:
            return p


def box_surface_points(rng, n, size, center):
    """Area-weighted uniform samples on the surface of an axis-aligned box (trimesh.sample stand-in,
    examples/detect_collision.py:36-46)."""
    sx, sy, sz = size
    areas = np.array([sy * sz, sy * sz, sx * sz, sx * sz, sx * sy, sx * sy])
    face = rng.choice(6, size=n, p=areas / areas.sum())
    uv = rng.random((n, 2)) - 0.5
    pts = np.zeros((n, 3))
    for f in range(6):
        m = face == f
        ax = f // 2
        sign = 1.0 if f % 2 == 0 else -1.0
        o = [a for a in range(3) if a != ax]
        pts[m, ax] = sign * 0.5 * size[ax]
        pts[m, o[0]] = uv[m, 0] * size[o[0]]
        pts[m, o[1]] = uv[m, 1] * size[o[1]]
    return pts + np.asarray(center)[None, :]


def add_c3_primitives(world, n_each=32, seed=3, keep=None):
    """n_each spheres, boxes, capsules, cylinders in the half-shell 0.3<=|p|<=1, z>=0."""
    from mplib_b200.pymp import fcl

    rng = np.random.Generator(np.random.PCG64(seed))
    k = 0
    for kind in ("sphere", "box", "capsule", "cylinder"):
        for _ in range(n_each):
            if kind == "sphere":
                g = fcl.Sphere(rng.uniform(0.02, 0.10))
            elif kind == "box":
                g = fcl.Box(rng.uniform(0.04, 0.20, 3))
            elif kind == "capsule":
                g = fcl.Capsule(rng.uniform(0.02, 0.05), rng.uniform(0.05, 0.30))
            else:
                g = fcl.Cylinder(rng.uniform(0.02, 0.05), rng.uniform(0.05, 0.30))
            pos, rot = _shell_point(rng), _rand_rot(rng)
            if keep is None or keep(kind, k, pos):
                world.add_normal_object(f"{kind}_{k}", fcl.CollisionObject(g, pos, rot))
            k += 1


def attach_c3_box(world):
    """Grasped box on panda_hand (examples/collision_avoidance.py:88-90); touch links = hand + fingers."""
    from mplib_b200.pymp import fcl

    art = world.get_articulation("robot")
    link_id = art.get_user_link_names().index("panda_hand")
    world.attach_object("grasped", fcl.Box([0.04, 0.04, 0.12]), "robot", link_id, [0, 0, 0.14, 1, 0, 0, 0],
                        ["panda_hand", "panda_leftfinger", "panda_rightfinger"])


def c3_world(n_each=32, mesh=True, cloud_points=100000, primitives=True, attach=True):
    import os

    from mplib_b200.pymp import fcl

    w = make_panda_world(convex=True)
    if primitives:
        add_c3_primitives(w, n_each)
    if attach:
        attach_c3_box(w)
    if mesh:
        m = fcl.load_mesh_as_BVH(os.path.join(ROOT, "data/panda/franka_description/meshes/collision/link0.stl"))
        w.add_normal_object("mesh_obstacle", fcl.CollisionObject(m, [0.6, 0.3, 0.2], [1, 0, 0, 0]))
    if cloud_points:
        rng = np.random.Generator(np.random.PCG64(4))
        n1 = cloud_points // 2
        pts = np.concatenate([box_surface_points(rng, n1, (0.8, 0.8, 0.05), (0.56, 0, -0.025)),
                              box_surface_points(rng, cloud_points - n1, (0.05, 1.0, 0.8), (0.9, 0, 0.4))])
        w.add_point_cloud("cloud", pts, 0.01)
    return w


def c3(batch=0, **kw):
    """configs[2]: Panda + grasped box vs 128 primitives + 1 mesh + 100k-point octree, ACM-masked."""
    n = batch or (1 << 22)
    w = c3_world(**kw)
    desc = compile_desc(w)
    return {
        "world": w, "desc": desc, "batch": n,
        "make_q": lambda k, m: random_q(m, 2 + 7919 * k).astype(np.float32),
        "config": {"workload": "configs[2]: Panda + grasped box (AttachedBody) vs 128 primitives + 1 mesh + 100k-point "
                               f"octree, ACM-masked, {n} random configurations per GPU per step (seed 2)",
                   "states_per_gpu": n, "state_dim": desc.state_dim, "pairs": desc.n_pairs,
                   "outputs": "verdict bit + pair mask per configuration",
                   "l2_policy": "rotating pool of input batches > 126 MB"},
        "flop_model": _flop_model(desc),
        "hit_cap_per_state": 12,  # ~9.5 colliding pairs per configuration in this scene (obstacles fill the reach shell)
    }


def c4_world(cloud_points=50000):
    """SURVEY 8(d) C4 scene: the C3 scene restricted to its first 16 boxes + the table cloud (grasped box kept)."""
    w = make_panda_world(convex=True)
    # SURVEY's half-shell (0.3 <= |p| <= 1) reaches the robot's own base column: a box there intersects link1/link2 in
    # every configuration and no collision-free start/goal exists.  Planning scenes therefore skip boxes whose centre
    # lies within 0.30 m of the base axis segment x = y = 0, 0 <= z <= 0.65, and take the first 16 that remain.
    taken = []

    def keep(kind, k, pos):
        if kind != "box" or len(taken) >= 16:
            return False
        d = np.linalg.norm(pos - np.array([0.0, 0.0, np.clip(pos[2], 0.0, 0.65)]))
        if d < 0.30:
            return False
        taken.append(k)
        return True

    add_c3_primitives(w, 32, keep=keep)
    attach_c3_box(w)
    if cloud_points:
        rng = np.random.Generator(np.random.PCG64(4))
        w.add_point_cloud("cloud", box_surface_points(rng, cloud_points, (0.8, 0.8, 0.05), (0.56, 0, -0.025)), 0.01)
    return w


def free_states(world, n, seed, checker):
    """Rejection-sample n collision-free states; ``checker(states) -> colliding bool[n]``."""
    out, k = [], 0
    while sum(len(x) for x in out) < n:
        q = random_q(4 * n, seed + 7919 * k)
        out.append(q[~checker(q)])
        k += 1
    return np.concatenate(out)[:n]


def prefixed_panda_files(prefix, xyz, yaw):
    """URDF/SRDF of a Panda whose link and joint names carry ``prefix`` and whose base sits at ``xyz`` with ``yaw``:
    the reference has no base-pose API and the ACM is keyed by link name only, so the two robots of
    configs[4] are separate URDFs with an extra fixed root joint.  Files are generated next to panda.urdf so that the
    relative mesh paths keep resolving."""
    import os
    import re


    d = os.path.dirname(URDF)
    urdf_out = os.path.join(d, f"_gen_{prefix}panda.urdf")
    srdf_out = os.path.join(d, f"_gen_{prefix}panda.srdf")
    u = open(URDF).read()
    u = re.sub(r'(<(?:link|joint)\s+name=")', r"\g<1>" + prefix, u)
    u = re.sub(r'(<(?:parent|child)\s+link=")', r"\g<1>" + prefix, u)
    u = re.sub(r'(<mimic\s+joint=")', r"\g<1>" + prefix, u)
    root = (f'  <link name="{prefix}base"/>\n  <joint name="{prefix}base_joint" type="fixed">\n'
            f'    <origin rpy="0 0 {yaw!r}" xyz="{xyz[0]!r} {xyz[1]!r} {xyz[2]!r}"/>\n'
            f'    <parent link="{prefix}base"/>\n    <child link="{prefix}panda_link0"/>\n  </joint>\n')
    u = u.replace("</robot>", root + "</robot>")
    sr = open(SRDF).read()
    sr = re.sub(r'((?:link1|link2|link|parent_link|child_link|joint)\s*=\s*")', r"\g<1>" + prefix, sr)
    for path, txt in ((urdf_out, u), (srdf_out, sr)):
        if not os.path.exists(path) or open(path).read() != txt:
            tmp = f"{path}.{os.getpid()}.tmp"  # written aside and renamed: the ranks of a multi-GPU run generate it concurrently
            with open(tmp, "w") as f:
                f.write(txt)
            os.replace(tmp, path)
    return urdf_out, srdf_out


def c5_world():
    """configs[4]: two Pandas (a_ at y=-0.35 yaw 0, b_ at y=+0.35 yaw pi), both planned, cross-robot ACM entries."""
    from mplib_b200.pymp import ArticulatedModel, PlanningWorld

    arts = []
    for prefix, xyz, yaw in (("a_", (0.0, -0.35, 0.0), 0.0), ("b_", (0.0, 0.35, 0.0), float(np.pi))):
        urdf, srdf = prefixed_panda_files(prefix, xyz, yaw)
        art = ArticulatedModel(urdf, srdf, (0, 0, -9.81), [prefix + j for j in PANDA_JOINTS], [prefix + l for l in PANDA_LINKS],
                               verbose=False, convex=True)
        art.set_move_group(prefix + "panda_hand")
        arts.append(art)
    w = PlanningWorld(arts, ["robot_a", "robot_b"])
    acm = w.get_allowed_collision_matrix()
    acm.set_entry("a_panda_link0", "b_panda_link0", True)
    rng = np.random.Generator(np.random.PCG64(7))
    coll = [l for l in PANDA_LINKS if l != "panda_link8"]
    done = {("panda_link0", "panda_link0")}
    while len(done) < 9:  # 8 further random cross entries
        x, y = coll[rng.integers(len(coll))], coll[rng.integers(len(coll))]
        if (x, y) in done:
            continue
        done.add((x, y))
        acm.set_entry("a_" + x, "b_" + y, True)
    return w


def c5(batch=0):
    """configs[4]: Two-Panda PlanningWorld with cross-robot ACM, 14-dof states."""
    n = batch or (1 << 22)
    w = c5_world()
    desc = compile_desc(w)
    lo, hi = np.concatenate([PANDA_LO, PANDA_LO]), np.concatenate([PANDA_HI, PANDA_HI])
    return {
        "world": w, "desc": desc, "batch": n,
        "make_q": lambda k, m: random_q(m, 6 + 7919 * k, lo, hi).astype(np.float32),
        "config": {"workload": "configs[4]: two Pandas (bases at y=-0.35/+0.35, yaw 0/pi), 2x19 self + 11x11 cross pairs minus 9 "
                               f"cross-robot ACM entries, {n} random 14-dof configurations per GPU per step (seed 6)",
                   "states_per_gpu": n, "state_dim": desc.state_dim, "pairs": desc.n_pairs,
                   "outputs": "verdict bit + pair mask per configuration",
                   "l2_policy": "rotating pool of input batches > 126 MB"},
        "flop_model": _flop_model(desc),
    }


def c1_world(scene="B"):
    """configs[0]: Panda self+env check of 10k random qpos.  Scene "A" is examples/detect_collision.py:36-46 (10 000 points
    on the surface of a 2 x 2 x 0.1 m box centred (0, 0, -0.1), add_point_cloud at resolution 1e-3; trimesh's sampler is
    restated as an area-weighted face sampler, seed 100); scene "B" is the "box/sphere" reading: Box(2, 2, 0.1) at
    (0, 0, -0.1) + Sphere(0.1) at (0.5, 0, 0.3)."""
    from mplib_b200.pymp import fcl

    w = make_panda_world(convex=True)
    if scene == "A":
        rng = np.random.Generator(np.random.PCG64(100))
        w.add_point_cloud("floor", box_surface_points(rng, 10000, (2.0, 2.0, 0.1), (0.0, 0.0, -0.1)), 1e-3)
    else:
        w.add_normal_object("floor_box", fcl.CollisionObject(fcl.Box([2.0, 2.0, 0.1]), [0.0, 0.0, -0.1], [1, 0, 0, 0]))
        w.add_normal_object("ball", fcl.CollisionObject(fcl.Sphere(0.1), [0.5, 0.0, 0.3], [1, 0, 0, 0]))
    return w


def c1_q(n=10000):
    return random_q(n, 0)


def c4(batch=0):
    """configs[3]: RRTConnect plan_qpos queries in the C4 scene.  ``batch`` = planning queries per GPU per step."""
    n = batch or 8192
    w = c4_world()
    desc = compile_desc(w)
    return {
        "world": w, "desc": desc, "batch": n,
        "config": {"workload": "configs[3]: RRTConnect plan_qpos, C4 scene (Panda + grasped box vs 16 boxes + 50k-point table "
                               f"octree), {n} independent (start, goal) queries per GPU per step (seed 5) through concurrent search slots, range 0.1, validity "
                               "resolution 1 %, one goal per query, every state of every edge checked (DiscreteMotionValidator)",
                   "queries_per_gpu": n, "state_dim": desc.state_dim, "pairs": desc.n_pairs,
                   "outputs": "status + waypoint matrix per query",
                   "l2_policy": "each round rewrites the tree/edge/work-queue buffers (> L2 at this batch); queries differ per step"},
        "flop_model": _flop_model(desc),
    }


def c4_queries(n, seed, checker):
    """n (start, goal) pairs, both collision-free under ``checker(states) -> colliding bool``."""
    pts = free_states(None, 2 * n, seed, checker)
    return pts[:n], pts[n:]


def get(name, batch=0):
    return {"c2": c2, "c3": c3, "c4": c4, "c5": c5}[name](batch)
