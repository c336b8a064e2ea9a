"""ctypes view of oracle/_ref/liboracle_ref*.so — the UNMODIFIED reference compiled in place by
oracle/Makefile (see oracle/ref_harness.cc).  TEST INFRASTRUCTURE ONLY."""
from __future__ import annotations

import ctypes as C
from pathlib import Path

import numpy as np

ORACLE_DIR = Path(__file__).resolve().parent


def available(fast: bool = False) -> bool:
    return (ORACLE_DIR / "_ref" / ("liboracle_ref_fast.so" if fast else "liboracle_ref.so")).exists()


_LIBS = {}


def _load(fast: bool):
    if fast in _LIBS:
        return _LIBS[fast]
    p = ORACLE_DIR / "_ref" / ("liboracle_ref_fast.so" if fast else "liboracle_ref.so")
    lib = C.CDLL(str(p))
    lib.ora_scene_create.restype = C.c_void_p
    lib.ora_scene_create.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_char_p]
    lib.ora_scene_free.argtypes = [C.c_void_p]
    lib.ora_intersect.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_int]
    lib.ora_intersect_world.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_int]
    lib.ora_intersect_world.restype = C.c_double
    lib.ora_intersect_records.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_int]
    lib.ora_intersect_records.restype = None
    lib.ora_bvh_count.argtypes = [C.c_void_p]
    lib.ora_bvh_size.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
    lib.ora_bvh_dump.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
    lib.ora_camera_dims.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    lib.ora_camera_rays.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_uint, C.c_void_p]
    lib.ora_render_a.restype = C.c_double
    lib.ora_render_a.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_uint, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                 C.c_void_p, C.c_void_p, C.c_int]
    lib.ora_render_b.restype = C.c_double
    lib.ora_render_b.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_uint, C.c_int, C.c_void_p, C.c_void_p, C.c_int]
    lib.ora_pixel_state_run.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.c_double, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.ora_pixel_state_run.restype = None
    lib.ora_texture_value.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p]
    _LIBS[fast] = lib
    return lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class RefScene:
    """The reference's own scene objects (geom::Bvh etc.) built from the product's export stream."""

    def __init__(self, prims, mats, texs, image_paths: str = "", fast: bool = False):
        self.lib = _load(fast)
        self._keep = [np.ascontiguousarray(prims, np.float64), np.ascontiguousarray(mats, np.float64), np.ascontiguousarray(texs, np.float64)]
        p, m, t = self._keep
        self.h = self.lib.ora_scene_create(_p(p), p.shape[0], _p(m), m.shape[0], _p(t), t.shape[0], image_paths.encode())
        if not self.h:
            raise RuntimeError("oracle: could not build the reference scene")

    def close(self):
        if self.h:
            self.lib.ora_scene_free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def intersect(self, rays: np.ndarray, threads: int = 0):
        """(t[n] float64, gid[n] int32) via the id-recovering restatement of Bvh::Hit."""
        r = np.ascontiguousarray(rays).view(np.float32).reshape(-1, 8)
        t = np.zeros(r.shape[0], np.float64)
        g = np.zeros(r.shape[0], np.int32)
        self.lib.ora_intersect(self.h, _p(r), r.shape[0], _p(t), _p(g), threads)
        return t, g

    def intersect_world(self, rays: np.ndarray, threads: int = 0):
        """(t, hit, seconds) via the reference's own scene::Scene::Hit -> Bvh::Hit."""
        r = np.ascontiguousarray(rays).view(np.float32).reshape(-1, 8)
        t = np.zeros(r.shape[0], np.float64)
        h = np.zeros(r.shape[0], np.uint8)
        secs = self.lib.ora_intersect_world(self.h, _p(r), r.shape[0], _p(t), _p(h), threads)
        return t, h.astype(bool), secs

    def intersect_records(self, rays: np.ndarray, threads: int = 0) -> np.ndarray:
        """The reference's HitRecord per ray over (0.001f, inf): [n, 13] = hit, t, p, normal, front_face, u, v, id, has_mat."""
        r = np.ascontiguousarray(rays).view(np.float32).reshape(-1, 8)
        out = np.zeros((r.shape[0], 13), np.float64)
        self.lib.ora_intersect_records(self.h, _p(r), r.shape[0], _p(out), threads)
        return out

    def bvh_count(self) -> int:
        return self.lib.ora_bvh_count(self.h)

    def bvh_dump(self, k: int):
        nn, npr = C.c_size_t(), C.c_size_t()
        self.lib.ora_bvh_size(self.h, k, C.byref(nn), C.byref(npr))
        nodes = np.zeros((nn.value, 9), np.float64)
        idx = np.zeros(npr.value, np.int32)
        self.lib.ora_bvh_dump(self.h, k, _p(nodes), _p(idx))
        return nodes, idx

    def render_a(self, cfg16: np.ndarray, spp: int, max_depth: int, seed: int = 1, threads: int = 0, y0: int = 0, y1: int = -1,
                 want_m2: bool = True, depth_camera: bool = False):
        """Path A (MegaKernel/DefaultSampler/GetPixel): mean[h,w,3], m2[h,w,3], seconds, rays."""
        cfg = np.ascontiguousarray(cfg16, np.float64)
        w, h = C.c_int(), C.c_int()
        self.lib.ora_camera_dims(_p(cfg), C.byref(w), C.byref(h))
        mean = np.zeros((h.value, w.value, 3), np.float64)
        m2 = np.zeros((h.value, w.value, 3), np.float64) if want_m2 else None
        rays = C.c_uint64()
        secs = self.lib.ora_render_a(self.h, _p(cfg), spp, max_depth, seed, threads, y0, y1, _p(mean), _p(m2), C.byref(rays), int(depth_camera))
        return mean, m2, secs, rays.value

    def render_b(self, cfg16: np.ndarray, max_depth: int, max_samples: int, seed: int = 1, threads: int = 0):
        """Path B as shipped (WavefrontRenderer::Render over CPURayIntegrator, adaptive stop 0.05 / 16):
        rgb8[h,w,3] parsed from its P3 output, active[s] = pixels still sampling in pass s, seconds."""
        cfg = np.ascontiguousarray(cfg16, np.float64)
        w, h = C.c_int(), C.c_int()
        self.lib.ora_camera_dims(_p(cfg), C.byref(w), C.byref(h))
        rgb8 = np.zeros((h.value, w.value, 3), np.uint8)
        active = np.zeros(max_samples, np.int32)
        secs = self.lib.ora_render_b(self.h, _p(cfg), max_depth, max_samples, seed, threads, _p(rgb8), _p(active), max_samples)
        if secs < 0:
            raise RuntimeError("oracle: could not parse the reference's P3 output")
        return rgb8, active, secs

    def texture_value(self, tex_id: int, uvp: np.ndarray) -> np.ndarray:
        q = np.ascontiguousarray(uvp, np.float64).reshape(-1, 5)
        out = np.zeros((q.shape[0], 3), np.float64)
        self.lib.ora_texture_value(self.h, tex_id, _p(q), q.shape[0], _p(out))
        return out


def camera_rays(cfg16, ij, seed: int = 1, fast: bool = False) -> np.ndarray:
    lib = _load(fast)
    cfg = np.ascontiguousarray(cfg16, np.float64)
    ij = np.ascontiguousarray(ij, np.int32).reshape(-1, 2)
    out = np.zeros((ij.shape[0], 6), np.float64)
    lib.ora_camera_rays(_p(cfg), _p(ij), ij.shape[0], seed, _p(out))
    return out


def pixel_state_run(samples, rel_threshold: float = 0.05, min_spp: int = 16, fast: bool = False):
    """The reference's own RecordSample / IsConverged on sample sequences [n, len, 3]: (stop_n, sum, m2)."""
    lib = _load(fast)
    a = np.ascontiguousarray(samples, np.float64)
    n, ln, _ = a.shape
    stop = np.zeros(n, np.int32)
    s, m2 = np.zeros((n, 3), np.float64), np.zeros((n, 3), np.float64)
    lib.ora_pixel_state_run(_p(a), n, ln, rel_threshold, min_spp, _p(stop), _p(s), _p(m2))
    return stop, s, m2


def num_threads(fast: bool = False) -> int:
    return _load(fast).ora_num_threads()
