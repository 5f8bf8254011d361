"""CPU tests of the drop-in boundary and the host logic: libmpb.so loads and exports every symbol
include/mpb.h declares, fails loudly without a GPU (no CPU fallback), and the host mirror of the reference
classes (ACM, attach/detach, pair enumeration order, error types) follows src/planning_world.cpp."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from tests.common import PANDA_JOINTS, PANDA_LINKS, ROOT, SRDF, URDF, compile_desc, has_gpu, make_panda, make_panda_world


def _declared_symbols():
    txt = open(os.path.join(ROOT, "include", "mpb.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(mpb_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol():
    from mplib_b200 import _lib

    lib = _lib.load()
    declared = _declared_symbols()
    assert len(declared) >= 10
    for name in declared:
        assert hasattr(lib, name), f"libmpb.so does not export {name}"
    assert sorted(_lib.SYMBOLS) == declared, "mplib_b200/_lib.py SYMBOLS is out of sync with include/mpb.h"


def test_struct_mirror_matches_header():
    """Field order of the ctypes mirror == field order of mpb_scene_desc / mpb_stats in the header."""
    from mplib_b200.scene import MpbSceneDesc, MpbStats

    txt = open(os.path.join(ROOT, "include", "mpb.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)

    def fields(struct_name):
        body = re.search(r"typedef struct %s \{(.*?)\} %s;" % (struct_name, struct_name), txt, flags=re.S).group(1)
        return [re.search(r"(\w+)\s*$", ln.strip().rstrip(";")).group(1) for ln in body.split("\n") if ln.strip().endswith(";")]

    assert fields("mpb_scene_desc") == [f for f, _ in MpbSceneDesc._fields_]
    assert fields("mpb_stats") == [f for f, _ in MpbStats._fields_]


@pytest.mark.skipif(has_gpu(), reason="this test asserts the no-GPU failure mode")
def test_no_gpu_means_loud_failure(panda_world):
    """There is no CPU compute path behind the C-ABI: without a device every query raises."""
    from mplib_b200 import _lib

    assert _lib.load().mpb_device_count() == 0
    with pytest.raises(_lib.MpbError, match="no CUDA device"):
        _lib.DeviceScene(compile_desc(panda_world))
    with pytest.raises(_lib.MpbError):
        panda_world.collide()
    v = C.c_float(0)
    assert _lib.load().mpb_fp32_peak(0, C.byref(v)) != 0


def test_product_does_not_import_oracle():
    """The oracle is test infrastructure: nothing under mplib_b200/ may reference it."""
    for dp, _, files in os.walk(os.path.join(ROOT, "mplib_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                src = open(os.path.join(dp, f), errors="replace").read()
                assert not re.search(r"^\s*(from|import)\s+\.*oracle|import_module\(.oracle|liboracle|#include.*oracle", src,
                                     flags=re.M), f"{f} imports / links the oracle"


# ------------------------------------------------------------------------------------------- ACM
def test_acm_default_and_entry_rules():
    """src/collision_matrix.cpp:147-164, 194-198."""
    from mplib_b200.pymp import AllowedCollision, AllowedCollisionMatrix

    acm = AllowedCollisionMatrix()
    assert acm.get_allowed_collision("a", "b") is None and acm.admits("a", "b")
    acm.set_entry("a", "b", True)
    assert acm.get_allowed_collision("a", "b") == AllowedCollision.ALWAYS and not acm.admits("b", "a")
    acm.set_entry("a", "b", False)
    assert acm.get_allowed_collision("b", "a") == AllowedCollision.NEVER and acm.admits("a", "b")
    acm.remove_entry("a", "b")
    assert acm.get_allowed_collision("a", "b") is None
    # defaults: one default wins; both: NEVER > CONDITIONAL > ALWAYS; an entry beats defaults
    acm.set_default_entry("x", True)
    assert acm.get_allowed_collision("x", "y") == AllowedCollision.ALWAYS
    acm.set_default_entry("y", False)
    assert acm.get_allowed_collision("x", "y") == AllowedCollision.NEVER
    acm.set_entry("x", "y", True)
    assert acm.get_allowed_collision("x", "y") == AllowedCollision.ALWAYS
    v0 = acm.version
    acm.clear()
    assert acm.version > v0 and acm.get_allowed_collision("x", "y") is None


def test_world_pair_enumeration_c3():
    """Pair sections and counts of SURVEY F9/F10 for the C3 scene: 19 self + 11 link x attached (3 touch links
    masked by the ACM) + (11 links + 1 attached) x 130 scene objects."""
    from tests.workloads import c3_world

    w = c3_world(cloud_points=2000)
    desc = compile_desc(w)
    kinds = {}
    for k in desc.pair_kinds:
        kinds[k] = kinds.get(k, 0) + 1
    assert kinds == {"self": 19, "self_attach": 8, "articulation_sceneobject": 11 * 130, "attach_sceneobject": 130}
    assert desc.n_pairs == 1587
    # attached body comes first in collide(attached, link) (src/planning_world.cpp:335)
    k = desc.pair_kinds.index("self_attach")
    assert desc.object_names[desc.pair_a[k]] == "grasped"
    # detaching removes the ACM entries and the attached pairs
    w.detach_object("grasped", also_remove=True)
    d2 = compile_desc(w)
    assert d2.n_pairs == 19 + 11 * 130


def test_world_error_types():
    """Error convention of SURVEY 8b: ASSERT -> RuntimeError, .at() -> IndexError, invalid_argument -> ValueError."""
    from mplib_b200.pymp import ArticulatedModel, PlanningWorld, fcl

    w = make_panda_world()
    with pytest.raises(RuntimeError):
        w.set_qpos_all(np.zeros(6))
    with pytest.raises(RuntimeError):
        w.set_qpos_all(np.zeros(8))
    with pytest.raises(IndexError):
        w.set_qpos("nope", np.zeros(7))
    with pytest.raises(IndexError):
        w.attach_object("ghost", "robot", 9, [0, 0, 0, 1, 0, 0, 0], [])
    with pytest.raises(ValueError):
        ArticulatedModel(URDF, SRDF, (0, 0, -9.81), PANDA_JOINTS, PANDA_LINKS[:5], verbose=False, convex=True)
    with pytest.raises(RuntimeError):
        PlanningWorld([make_panda()], ["a", "b"])
    art = w.get_articulation("robot")
    with pytest.raises(RuntimeError):
        art.set_qpos(np.zeros(3))
    assert art.get_qpos_dim() == 7 and art.get_move_group_joint_indices() == list(range(7))
    # set_qpos_all writes only the move group; fingers keep their value (src/articulated_model.cpp:108-113)
    art.set_qpos(np.arange(9) * 0.01, full=True)
    w.set_qpos_all(np.ones(7) * 0.5)
    assert np.allclose(art.get_qpos(), [0.5] * 7 + [0.07, 0.08])
    assert isinstance(fcl.Box([1, 2, 3]).side, np.ndarray)


def test_two_articulations_state_split_in_name_order():
    """PlanningWorldTpl::setQposAll splits the state over planned articulations in std::map (name) order
    (src/planning_world.cpp:249-262)."""
    from mplib_b200.pymp import PlanningWorld

    a, b = make_panda(), make_panda()
    w = PlanningWorld([a, b], ["zeta", "alpha"])
    w.set_qpos_all(np.concatenate([np.full(7, 0.1), np.full(7, 0.2)]))
    assert np.allclose(b.get_qpos()[:7], 0.1) and np.allclose(a.get_qpos()[:7], 0.2)
    desc = compile_desc(w)
    assert desc.state_dim == 14 and desc.n_joints == 18 and desc.n_objects == 22
    kinds = {}
    for k in desc.pair_kinds:
        kinds[k] = kinds.get(k, 0) + 1
    assert kinds == {"self": 38, "self_articulation": 121}


def test_native_host_module_builds_and_fails_loudly_without_gpu():
    """The pybind11 host module (csrc/pymp_native.cpp) is what the per-call path runs through; it must import, expose the
    scene class, and -- like the C-ABI under it -- refuse to work without a CUDA device instead of falling back."""
    from mplib_b200 import _lib
    from tests.common import compile_desc, make_panda_world

    assert _lib._native is not None, "mplib_b200/_native*.so is missing: run `make`"
    for name in ("collide", "collide_hits", "fk", "distance", "poll", "set_q_base", "set_pairs", "update_objects", "close", "handle"):
        assert hasattr(_lib._native.Scene, name)
    if _lib.load().mpb_device_count() == 0:
        with pytest.raises(_lib.MpbError, match="no CUDA device"):
            _lib.DeviceScene(compile_desc(make_panda_world()), 0)
