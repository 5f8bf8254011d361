"""Shared builders for the tests: the Panda world of the reference's own tests
(tests/test_basic.py:5-36) and seeded random configurations; defined in mplib_b200/workloads.py."""
from mplib_b200.workloads import (PANDA_HI, PANDA_JOINTS, PANDA_LINKS, PANDA_LO, ROOT, SRDF, URDF, compile_desc,  # noqa: F401
                                  make_panda, make_panda_world, random_q)


def has_gpu():
    try:
        from mplib_b200 import _lib

        return _lib.load().mpb_device_count() > 0
    except Exception:
        return False
