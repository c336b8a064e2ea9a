"""3360-ray-tracer_b200 — B200-native path-tracing hot path behind the reference's scene /
renderer API.  The product is librtb200.so (CUDA kernels + C ABI, include/rtb.h) and the host
C++ scene API (host/rt/...); this Python package is the ctypes plumbing used by tests and
bench.py.  The directory name is not a valid Python identifier: load it with
`__graft_entry__.load_package()` (registers it as module `rt_b200`).
"""
from .binding import *  # noqa: F401,F403
from . import binding  # noqa: F401
from . import multigpu  # noqa: F401,E402
