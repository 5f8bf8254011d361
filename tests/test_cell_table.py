"""Support cell table of a convex hull (scene compiler, host side): the device scans only the candidates of the cube-map
cell of the support direction, so the arg-max vertex of ANY direction has to be among them.  Checked on the Panda's hulls
through the diagnostics entry point mpb_debug_cell_table; needs no GPU."""
import numpy as np

from mplib_b200 import _lib
from mplib_b200.workloads import compile_desc, make_panda_world

G = 8


def cube_cell(d):
    """mplib_b200/csrc/mpb_device.cuh: mpb_cube_cell, vectorised."""
    a = np.abs(d)
    x, y, z = d[:, 0], d[:, 1], d[:, 2]
    fx = (a[:, 0] >= a[:, 1]) & (a[:, 0] >= a[:, 2])
    fy = ~fx & (a[:, 1] >= a[:, 2])
    face = np.where(fx, np.where(x >= 0, 0, 1), np.where(fy, np.where(y >= 0, 2, 3), np.where(z >= 0, 4, 5)))
    m = np.where(fx, a[:, 0], np.where(fy, a[:, 1], a[:, 2]))
    u = np.where(fx, y, x)
    v = np.where(fx | fy, z, y)
    iu = np.clip(((u / m + 1.0) * (0.5 * G)).astype(np.int64), 0, G - 1)
    iv = np.clip(((v / m + 1.0) * (0.5 * G)).astype(np.int64), 0, G - 1)
    return (face * G + iv) * G + iu


def test_argmax_vertex_is_a_candidate_of_its_cell():
    desc = compile_desc(make_panda_world())
    off = np.asarray(desc.hull_vert_off)
    V = np.asarray(desc.hull_verts).reshape(-1, 3)
    rng = np.random.default_rng(5)
    d = rng.normal(size=(200000, 3)).astype(np.float32).astype(np.float64)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    cells = cube_cell(d.astype(np.float32))
    sizes = []
    for h in range(len(off) - 1):
        v = V[off[h]:off[h + 1]]
        if len(v) > 255:
            continue
        first, cand = _lib.debug_cell_table(v)
        assert first[0] == 0 and first[-1] == len(cand) and np.all(np.diff(first.astype(np.int64)) >= 1)
        sizes.append(np.diff(first.astype(np.int64)).mean())
        best = (d @ v.T).argmax(axis=1)
        member = np.zeros((6 * G * G, len(v)), dtype=bool)
        for c in range(6 * G * G):
            member[c, cand[first[c]:first[c + 1]]] = True
        ok = member[cells, best]
        # a direction whose arg-max is not listed must be a numerical tie: the best listed vertex is as good to 1e-6 relative
        if not ok.all():
            bad = np.nonzero(~ok)[0]
            dots = d[bad] @ v.T
            listed = np.where(member[cells[bad]], dots, -np.inf).max(axis=1)
            assert np.all(dots.max(axis=1) - listed <= 1e-6 * np.abs(v).max()), (h, len(bad))
    assert sizes and max(sizes) < 8, sizes  # ~3 candidates per cell for the Panda's hulls
    print("mean candidates per cell, per hull:", np.round(sizes, 2))
