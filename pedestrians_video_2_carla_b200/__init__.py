"""B200-native FK + projection + loc_2d_3d hot path of pedestrians-video-2-carla.

Host code is PyTorch plumbing (device memory, streams, autograd graph); every op is a
hand-written sm_100a kernel behind the C ABI of include/p2c_b200.h (libp2c_b200.so, built
in-tree by `__graft_entry__.build()`).  There is no CPU fallback.
"""
from . import _build, _lib, ops  # noqa: F401

__version__ = '0.1'
