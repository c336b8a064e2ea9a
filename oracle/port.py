"""ctypes view of oracle/liboracle_port.so — the plain-C restatement of the reference's closest-hit
path (oracle/port.c).  TEST INFRASTRUCTURE ONLY."""
from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

ORACLE_DIR = Path(__file__).resolve().parent
_LIB = None


def _load():
    global _LIB
    if _LIB is not None:
        return _LIB
    so = ORACLE_DIR / "liboracle_port.so"
    if not so.exists():
        subprocess.run(["make", "port"], cwd=ORACLE_DIR, check=True)
    lib = C.CDLL(str(so))
    lib.port_scene_create.restype = C.c_void_p
    lib.port_scene_create.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.port_scene_free.argtypes = [C.c_void_p]
    lib.port_prim_count.argtypes = [C.c_void_p]
    lib.port_intersect.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_int]
    lib.port_pixel_state_run.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.c_double, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.port_pixel_state_run.restype = None
    _LIB = lib
    return lib


def pixel_state_run(samples, rel_threshold: float = 0.05, min_spp: int = 16):
    """oracle/port.c restatement of RecordSample / IsConverged on sample sequences [n, len, 3]: (stop_n, sum, m2)."""
    lib = _load()
    a = np.ascontiguousarray(samples, np.float64)
    n, ln, _ = a.shape
    stop = np.zeros(n, np.int32)
    s, m2 = np.zeros((n, 3), np.float64), np.zeros((n, 3), np.float64)
    lib.port_pixel_state_run(a.ctypes.data, n, ln, rel_threshold, min_spp, stop.ctypes.data, s.ctypes.data, m2.ctypes.data)
    return stop, s, m2


class PortOracle:
    """Closest hit (t, gid) by the C restatement.  The BVH arrays are the product host builder's
    (pinned against the reference's geom::Bvh by tests/test_host_scene.py); all intersection
    arithmetic is oracle/port.c."""
    kind = "port"

    def __init__(self, host_scene):
        self.lib = _load()
        prims, _, _, _ = host_scene.export()
        prims = np.ascontiguousarray(prims, np.float64)
        trees = [host_scene.bvh_dump(k) for k in range(host_scene.bvh_count())]
        nodes = [np.ascontiguousarray(n, np.float64) for n, _ in trees]
        idx = [np.ascontiguousarray(i, np.int32) for _, i in trees]
        k = len(trees)
        node_ptrs = (C.c_void_p * max(k, 1))(*[n.ctypes.data for n in nodes])
        idx_ptrs = (C.c_void_p * max(k, 1))(*[i.ctypes.data for i in idx])
        n_nodes = (C.c_size_t * max(k, 1))(*[n.shape[0] for n in nodes])
        n_idx = (C.c_size_t * max(k, 1))(*[i.shape[0] for i in idx])
        self.h = self.lib.port_scene_create(prims.ctypes.data, prims.shape[0], k, node_ptrs, n_nodes, idx_ptrs, n_idx)
        if not self.h:
            raise RuntimeError("oracle port: could not build the scene")

    def close(self):
        if getattr(self, "h", None):
            self.lib.port_scene_free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def intersect(self, rays: np.ndarray, threads: int = 0):
        r = np.ascontiguousarray(rays).view(np.float32).reshape(-1, 8)
        t = np.zeros(r.shape[0], np.float64)
        g = np.zeros(r.shape[0], np.int32)
        self.lib.port_intersect(self.h, r.ctypes.data, r.shape[0], t.ctypes.data, g.ctypes.data, threads)
        return t, g

    def render_a(self, *a, **kw):
        raise NotImplementedError("oracle/port.c restates the closest-hit path only; image oracles need oracle/_ref (make ref)")
