"""Re-export of the synthetic workloads (mplib_b200/workloads.py) under the name the tests use."""
from mplib_b200.workloads import *  # noqa: F401,F403
from mplib_b200.workloads import _flop_model, _rand_rot, _shell_point  # noqa: F401
