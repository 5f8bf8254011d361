"""The oracle against the committed fixtures of tests/golden/ (see tests/golden/make_golden.py for provenance):
hand-derived FK answers and the 19-pair list, and regression vectors of the oracle itself on C1/C2/C5 inputs."""
import json
import os

import numpy as np
import pytest

from tests.common import compile_desc, make_panda_world

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_golden(name):
    z = np.load(os.path.join(GOLD, name))
    P = int(z["n_pairs"])
    q = z["q"]
    v = np.unpackbits(z["verdict"], bitorder="little")[: q.shape[0]].astype(bool)
    ph = np.unpackbits(z["pairs"], axis=1, bitorder="little")[:, :P].astype(bool)
    return q, v, ph, [str(s) for s in z["pair_names"]]


def pair_names(desc):
    return [f"{desc.pair_kinds[p]}|{desc.object_names[desc.pair_a[p]]}|{desc.object_names[desc.pair_b[p]]}" for p in range(desc.n_pairs)]


def _quat_R(q):
    from mplib_b200.urdf import quat_to_mat

    return quat_to_mat(*q)


def test_fk_known_answers_fixture(panda_world):
    from oracle import oracle as O

    desc = compile_desc(panda_world)
    cases = json.load(open(os.path.join(GOLD, "fk_known_answers.json")))["cases"]
    for c in cases:
        fk = O.fk_batch(desc, np.asarray(c["qpos"][:7], dtype=np.float64)[None, :])[0]
        p = {n: fk[i] for i, n in enumerate(desc.link_names)}
        for link, exp in c["links"].items():
            R = _quat_R(p[link][3:])
            if "p" in exp:
                assert np.allclose(p[link][:3], exp["p"], atol=2e-6), link
            if "R_diag" in exp:
                assert np.allclose(np.diag(R), exp["R_diag"], atol=2e-6), link
            if "R_row0" in exp:
                assert np.allclose(R[0], exp["R_row0"], atol=2e-6), link
            if "R_row2" in exp:
                assert np.allclose(R[2], exp["R_row2"], atol=2e-6), link


def test_pair_list_fixture(panda_world):
    fm = panda_world.get_articulation("robot").get_fcl_model()
    names = fm.collision_link_names
    got = [[names[a], names[b]] for a, b in fm.collision_pairs]
    exp = json.load(open(os.path.join(GOLD, "panda_pairs.json")))["pairs"]
    assert sorted(map(tuple, got)) == sorted(map(tuple, exp)) and len(got) == 19


@pytest.mark.parametrize("fixture,builder", [("c1_sceneA.npz", "c1A"), ("c1_sceneB.npz", "c1B"), ("c2_self_20k.npz", "c2"),
                                             ("c5_two_panda_4k.npz", "c5")])
def test_oracle_reproduces_fixture(fixture, builder):
    """The oracle is deterministic: a 2 000-state slice of every fixture must come back bit-exact (guards the
    restatement and the scene compiler against silent drift; the full files are the GPU tests' reference)."""
    from oracle import oracle as O
    from tests import workloads as WL

    if builder == "c1A":
        desc = compile_desc(WL.c1_world("A"))
    elif builder == "c1B":
        desc = compile_desc(WL.c1_world("B"))
    elif builder == "c2":
        desc = compile_desc(make_panda_world())
    else:
        desc = WL.c5(4096)["desc"]
    q, v, ph, names = load_golden(fixture)
    assert names == pair_names(desc)
    n = min(2000, q.shape[0])
    ov, oph = O.collide_batch(desc, q[:n].astype(np.float64), threads=8)
    assert np.array_equal(oph, ph[:n]) and np.array_equal(ov, v[:n])
    assert np.array_equal(v, ph.any(axis=1))


def test_c1_example_spot_checks():
    """examples/detect_collision.py:36-65: with the floor point cloud, the self-collision-free pose stays free and a pose
    reaching below the floor collides with it (collision_type 'articulation_sceneobject' rows of the fixture)."""
    q, v, ph, names = load_golden("c1_sceneA.npz")
    env_cols = [i for i, s in enumerate(names) if s.startswith("articulation_sceneobject|")]
    assert len(env_cols) == 11
    assert ph[:, env_cols].any() and ph[:, [i for i in range(len(names)) if i not in env_cols]].any()
    # link0 rests on z = 0, the floor top is z = -0.05: never in contact
    assert not ph[:, [i for i in env_cols if "|panda_link0|" in names[i]]].any()
