"""mplib_b200: B200-native batched state-validity engine behind MPlib's pymp API surface."""
from . import pymp  # noqa: F401

__version__ = "0.1.0"
