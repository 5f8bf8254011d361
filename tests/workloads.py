"""Synthetic workloads of SURVEY 8(d) / BASELINE.json configs (seeds and shapes fixed there)."""
import numpy as np

from tests.common import PANDA_HI, PANDA_LO, compile_desc, make_panda_world, random_q


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


def get(name, batch=0):
    return {"c2": c2}[name](batch)
