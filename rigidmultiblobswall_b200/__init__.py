"""B200-native blob mobility product (Mf) for RigidMultiblobsWall.

Only the hot path of the reference is here: the dense
O(N^2) wall-corrected RPY/Blake mobility product behind the unchanged
mobility.py signatures, plus (later rows) the rigid-body operators around it.
"""
from . import _lib  # noqa: F401
from . import mobility  # noqa: F401

__version__ = "0.1.2"   # = RMB_VERSION 105 (include/rmb_mobility.h)
