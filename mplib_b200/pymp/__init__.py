"""``mplib.pymp`` mirror: same submodule and class names as the reference extension module
(``python/pybind.cpp:16-27``), implemented over the C-ABI of libmpb.so."""
from . import fcl  # noqa: F401
from . import collision_matrix  # noqa: F401
from . import articulation  # noqa: F401
from . import planning_world  # noqa: F401
from . import ompl  # noqa: F401
from .articulation import ArticulatedModel, FCLModel, PinocchioModel  # noqa: F401
from .collision_matrix import AllowedCollision, AllowedCollisionMatrix  # noqa: F401
from .planning_world import AttachedBody, PlanningWorld, WorldCollisionResult, WorldDistanceResult  # noqa: F401
from .ompl import OMPLPlanner, set_global_seed  # noqa: F401

fcl.FCLModel = FCLModel  # the reference exposes FCLModel inside pymp.fcl (python/pybind_fcl.hpp:439-466)


class pinocchio:  # namespace shim: pymp.pinocchio.PinocchioModel
    PinocchioModel = PinocchioModel
