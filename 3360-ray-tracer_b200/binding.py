"""ctypes binding of librtb200.so (include/rtb.h + include/rth.h).

This is plumbing for tests and bench.py; the product is the C ABI itself.  There is no
Python or CPU implementation of the hot path behind these calls: if the shared library is
missing, importing fails; if no CUDA device is present, compute calls raise RtbError.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import numpy as np

PKG_DIR = Path(__file__).resolve().parent
REPO_ROOT = PKG_DIR.parent
ASSET_DIR = REPO_ROOT / "assets"

RTB_OK, RTB_EINVAL, RTB_ENODEVICE, RTB_ECUDA, RTB_ENOMEM = 0, -1, -2, -3, -4
PRIM_SPHERE, PRIM_TRIANGLE, PRIM_RECT, PRIM_SUBTREE = 0, 1, 2, 3
MODE_A, MODE_B = 0, 1
RENDER_SMEM_TOP, RENDER_KEEP_ACCUM, RENDER_SERIAL, RENDER_WAVEFRONT, RENDER_NO_DRAIN, RENDER_POOL_PATHS, RENDER_COST_ORDER = 1, 2, 4, 8, 16, 32, 64
RENDER_WIDE4 = 128  # the 4-wide tree (scenes built with RTB_WIDE4=1 in the environment): an experiment, DESIGN.md 3.9
ADAPT_PIXEL_STATE, ADAPT_LUMINANCE = 0, 1
STATIC_BATCH_MAX = 65536  # RTB_STATIC_BATCH_MAX: most rays of one rtb_intersect_batch_static call
EX_SPHERE, EX_TRIANGLE, EX_RECT, EX_BVH_BEGIN, EX_BVH_END, EX_LIST_BEGIN, EX_LIST_END = 0, 1, 2, 10, 11, 12, 13
MAT_LAMBERTIAN, MAT_METAL, MAT_DIELECTRIC, MAT_DIFFUSE_LIGHT = 0, 1, 2, 3
TEX_SOLID, TEX_CHECKER, TEX_IMAGE = 0, 1, 2


class RtbError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"rtb error {code}: {msg}")
        self.code = code


class RtbNode(C.Structure):
    _fields_ = [("lo", C.c_float * 3), ("a", C.c_uint32), ("hi", C.c_float * 3), ("b", C.c_uint32)]


class RtbSceneDesc(C.Structure):
    _fields_ = [
        ("abi_version", C.c_uint32), ("n_nodes", C.c_uint32), ("nodes", C.c_void_p),
        ("n_prim_refs", C.c_uint32), ("prim_refs", C.c_void_p),
        ("n_spheres", C.c_uint32), ("spheres", C.c_void_p), ("sphere_mat", C.c_void_p),
        ("n_triangles", C.c_uint32), ("triangles", C.c_void_p),
        ("n_rects", C.c_uint32), ("rects", C.c_void_p),
        ("sphere_gid", C.c_void_p), ("triangle_gid", C.c_void_p), ("rect_gid", C.c_void_p),
        ("sphere_rank", C.c_void_p), ("triangle_rank", C.c_void_p), ("rect_rank", C.c_void_p),
        ("n_materials", C.c_uint32), ("materials", C.c_void_p),
        ("n_textures", C.c_uint32), ("textures", C.c_void_p),
        ("n_images", C.c_uint32), ("images", C.c_void_p),
        ("n_top_nodes", C.c_uint32), ("max_depth", C.c_uint32), ("compiled", C.c_void_p), ("flags", C.c_uint32),
    ]


class RtbCamera(C.Structure):
    _fields_ = [
        ("width", C.c_int32), ("height", C.c_int32),
        ("center", C.c_float * 3), ("pixel00", C.c_float * 3), ("delta_u", C.c_float * 3),
        ("delta_v", C.c_float * 3), ("disk_u", C.c_float * 3), ("disk_v", C.c_float * 3),
        ("defocus_angle", C.c_float), ("depth_camera", C.c_int32),
    ]


class RtbRenderParams(C.Structure):
    _fields_ = [
        ("spp_begin", C.c_int32), ("spp_end", C.c_int32), ("max_depth", C.c_int32), ("mode", C.c_int32),
        ("seed", C.c_uint64), ("pool_size", C.c_uint32), ("flags", C.c_uint32),
    ]


class RtbAdaptive(C.Structure):
    _fields_ = [("rel_threshold", C.c_float), ("min_samples", C.c_int32), ("criterion", C.c_int32), ("pad", C.c_int32)]


class RtbStats(C.Structure):
    _fields_ = [
        ("samples", C.c_uint64), ("rays", C.c_uint64), ("node_visits", C.c_uint64), ("sphere_tests", C.c_uint64),
        ("triangle_tests", C.c_uint64), ("rect_tests", C.c_uint64), ("iterations", C.c_uint32), ("launches", C.c_uint32),
        ("ms_total", C.c_float), ("ms_generate", C.c_float), ("ms_extend", C.c_float), ("ms_shade", C.c_float),
        ("ms_drain", C.c_float), ("drain_paths", C.c_uint32),
    ]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class RthCameraConfig(C.Structure):
    _fields_ = [
        ("aspect_ratio", C.c_double), ("image_width", C.c_int32), ("samples_per_pixel", C.c_int32),
        ("max_depth", C.c_int32), ("pad", C.c_int32), ("vfov", C.c_double),
        ("lookfrom", C.c_double * 3), ("lookat", C.c_double * 3), ("vup", C.c_double * 3),
        ("defocus_angle", C.c_double), ("focus_dist", C.c_double),
    ]

    def as_oracle_cfg(self) -> np.ndarray:
        """16 doubles in the order oracle/ref_harness.cc:SetupCamera reads."""
        return np.array([self.aspect_ratio, self.image_width, self.samples_per_pixel, self.max_depth, self.vfov,
                         *self.lookfrom, *self.lookat, *self.vup, self.defocus_angle, self.focus_dist], dtype=np.float64)


RAY_DTYPE = np.dtype([("o", np.float32, 3), ("tmin", np.float32), ("d", np.float32, 3), ("tmax", np.float32)])
HIT_DTYPE = np.dtype([("t", np.float64), ("prim_id", np.int32), ("ref", np.uint32)])
assert RAY_DTYPE.itemsize == 32 and HIT_DTYPE.itemsize == 16

# Every symbol include/rtb.h and include/rth.h declare (checked by tests/test_abi.py).
RTB_SYMBOLS = [
    "rtb_last_error", "rtb_abi_version", "rtb_tree_compile", "rtb_tree_destroy", "rtb_tree_info", "rtb_device_count", "rtb_debug_check_failures", "rtb_debug_fast_div", "rtb_init", "rtb_shutdown", "rtb_scene_create",
    "rtb_scene_destroy", "rtb_intersect_batch", "rtb_intersect_batch_static", "rtb_intersect_batch_device", "rtb_render", "rtb_render_device",
    "rtb_render_adaptive", "rtb_render_adaptive_device", "rtb_resolve_counts", "rtb_resolve_counts_device", "rtb_pixel_state_fold", "rtb_render_multi",
    "rtb_resolve_device", "rtb_resolve", "rtb_measure_l2_gbs", "rtb_measure_fp32_tflops", "rtb_measure_fp64_tflops",
]
RTH_SYMBOLS = [
    "rth_last_error", "rth_set_tree_compiler", "rth_scene_build", "rth_scene_build_ex", "rth_scene_from_records", "rth_scene_desc", "rth_scene_export",
    "rth_scene_gid_map", "rth_scene_bvh_count", "rth_scene_bvh_size", "rth_scene_bvh_dump", "rth_scene_free",
    "rth_camera_init", "rth_camera_preset", "rth_encode_png",
]


def lib_path(counters: bool = False) -> Path:
    # RTB_LIB=<file name in this directory>: a tuning variant built by `make VARIANT=...` (tools/gpu_variants.py)
    if not counters and os.environ.get("RTB_LIB"):
        return PKG_DIR / os.environ["RTB_LIB"]
    return PKG_DIR / ("librtb200_counters.so" if counters else "librtb200.so")


def host_lib_path() -> Path:
    return PKG_DIR / "librth.so"


_LIBS: dict = {}
_HOST: list = []
# rth_set_tree_compiler's callback types (include/rth.h)
_COMPILE_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p)
_DESTROY_FN = C.CFUNCTYPE(None, C.c_void_p)
_ERROR_FN = C.CFUNCTYPE(C.c_char_p)


def load_host() -> C.CDLL:
    """librth.so alone: the host C++ scene API (scene builders, flattener, cameras, PNG).  Pure C++ - loading it does not map
    the CUDA library, which is what `bench.py --impl reference` and the oracle-only tests use."""
    if _HOST:
        return _HOST[0]
    p = host_lib_path()
    if not p.exists():
        raise ImportError(f"{p} is missing: run `python -c 'import __graft_entry__ as g; g.build()'`")
    lib = C.CDLL(str(p))
    lib.rth_last_error.restype = C.c_char_p
    lib.rth_set_tree_compiler.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    lib.rth_set_tree_compiler.restype = None
    lib.rth_scene_desc.restype = C.POINTER(RtbSceneDesc)
    lib.rth_scene_desc.argtypes = [C.c_void_p]
    lib.rth_scene_build.argtypes = [C.c_char_p, C.c_void_p, C.c_int, C.c_char_p, C.c_uint32, C.POINTER(C.c_void_p)]
    lib.rth_scene_build_ex.argtypes = [C.c_char_p, C.c_void_p, C.c_int, C.c_char_p, C.c_uint32, C.c_uint32, C.POINTER(C.c_void_p)]
    lib.rth_scene_from_records.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t,
                                           C.c_char_p, C.c_uint32, C.POINTER(C.c_void_p)]
    lib.rth_scene_export.argtypes = [C.c_void_p] + [C.c_void_p] * 7
    lib.rth_scene_gid_map.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    lib.rth_scene_bvh_count.argtypes = [C.c_void_p]
    lib.rth_scene_bvh_size.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
    lib.rth_scene_bvh_dump.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
    lib.rth_scene_free.argtypes = [C.c_void_p]
    lib.rth_scene_free.restype = None
    lib.rth_camera_init.argtypes = [C.POINTER(RthCameraConfig), C.c_int, C.POINTER(RtbCamera)]
    lib.rth_camera_preset.argtypes = [C.c_char_p, C.c_char_p, C.POINTER(RthCameraConfig)]
    lib.rth_encode_png.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]
    _HOST.append(lib)
    return lib


class _Product:
    """Both libraries behind one object: rth_* resolves in librth.so, rtb_* in librtb200.so (or a counters / variant build)."""

    def __init__(self, host, cuda):
        self._host, self._cuda = host, cuda

    def __getattr__(self, name):
        return getattr(self._host if name.startswith("rth_") else self._cuda, name)


def load(counters: bool = False) -> "_Product":
    """Load the CUDA extension (and the host library).  Fails loudly when it has not been built — there is no fallback."""
    key = bool(counters)
    if key in _LIBS:
        return _LIBS[key]
    p = lib_path(counters)
    if not p.exists():
        raise ImportError(f"{p} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` (nvcc, sm_100a). "
                          "There is no CPU or PyTorch fallback for the path-tracing hot path.")
    host = load_host()
    lib = C.CDLL(str(p))
    lib.rtb_last_error.restype = C.c_char_p
    lib.rtb_scene_create.argtypes = [C.POINTER(RtbSceneDesc), C.POINTER(C.c_void_p)]
    lib.rtb_scene_destroy.argtypes = [C.c_void_p]
    lib.rtb_intersect_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
    lib.rtb_intersect_batch_device.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p]
    lib.rtb_render.argtypes = [C.c_void_p, C.POINTER(RtbCamera), C.POINTER(RtbRenderParams), C.c_void_p, C.POINTER(RtbStats)]
    lib.rtb_render_device.argtypes = [C.c_void_p, C.POINTER(RtbCamera), C.POINTER(RtbRenderParams), C.c_void_p, C.POINTER(RtbStats)]
    lib.rtb_render_adaptive.argtypes = [C.c_void_p, C.POINTER(RtbCamera), C.POINTER(RtbRenderParams), C.POINTER(RtbAdaptive), C.c_void_p,
                                        C.c_void_p, C.c_void_p, C.POINTER(RtbStats)]
    lib.rtb_render_adaptive_device.argtypes = lib.rtb_render_adaptive.argtypes
    lib.rtb_render_multi.argtypes = [C.POINTER(RtbSceneDesc), C.c_int, C.c_void_p, C.POINTER(RtbCamera), C.POINTER(RtbRenderParams), C.c_void_p,
                                     C.POINTER(RtbStats)]
    lib.rtb_pixel_state_fold.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.POINTER(RtbAdaptive), C.c_void_p, C.c_void_p, C.c_void_p]
    lib.rtb_resolve_counts.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
    lib.rtb_resolve_counts_device.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p]
    lib.rtb_resolve.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.c_void_p]
    lib.rtb_resolve_device.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.c_void_p, C.c_void_p]
    lib.rtb_measure_l2_gbs.argtypes = [C.c_size_t, C.c_int, C.POINTER(C.c_float)]
    lib.rtb_measure_fp32_tflops.argtypes = [C.c_int, C.POINTER(C.c_float)]
    lib.rtb_measure_fp64_tflops.argtypes = [C.c_int, C.POINTER(C.c_float)]
    lib.rtb_init.argtypes = [C.c_int]
    lib.rtb_debug_check_failures.argtypes = [C.c_void_p]
    lib.rtb_debug_fast_div.argtypes = [C.c_uint32, C.c_uint32]
    lib.rtb_debug_fast_div.restype = C.c_uint32
    prod = _Product(host, lib)
    _LIBS[key] = prod
    return prod


def _use_compiler(lib):
    """Scenes built from now on are compiled (rtb_tree_compile: host work, once per scene) by THIS device library; a scene
    keeps the destroyer of the library that compiled it."""
    cuda = lib._cuda
    lib._host.rth_set_tree_compiler(C.cast(cuda.rtb_tree_compile, C.c_void_p), C.cast(cuda.rtb_tree_destroy, C.c_void_p),
                                    C.cast(cuda.rtb_last_error, C.c_void_p))


def _check(lib, rc: int, host: bool = False):
    if rc != RTB_OK:
        msg = (lib.rth_last_error() if host else lib.rtb_last_error()) or b""
        raise RtbError(rc, msg.decode(errors="replace"))


def _ptr(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


class HostScene:
    """A scene built by the host C++ API (rt::scene::Scene + rt::geom::Bvh) and flattened."""

    def __init__(self, handle, lib):
        self._h, self._lib = handle, lib

    @classmethod
    def build(cls, name: str, params=(), top_budget: int = 0, counters: bool = False, asset_dir=None, device_tree: bool = False,
              host_only: bool = False) -> "HostScene":
        """device_tree: no host SAH build; rtb_scene_create builds a linear BVH on the GPU (RTB_SCENE_DEVICE_BUILD).
        host_only: librth.so alone (the CUDA library is not loaded; the scene carries no compiled device tree) - for feeding
        the oracle from a process that must not touch the product's kernels."""
        if host_only:
            lib = load_host()
            lib.rth_set_tree_compiler(None, None, None)
        else:
            lib = load(counters)
            _use_compiler(lib)
        p = np.asarray(params, dtype=np.float64)
        h = C.c_void_p()
        rc = lib.rth_scene_build_ex(name.encode(), _ptr(p) if p.size else None, int(p.size),
                                    str(asset_dir or ASSET_DIR).encode(), top_budget, 1 if device_tree else 0, C.byref(h))
        _check(lib, rc, host=True)
        return cls(h, lib)

    @classmethod
    def from_records(cls, prims, mats, texs, image_paths: str = "", top_budget: int = 0, counters: bool = False) -> "HostScene":
        lib = load(counters)
        _use_compiler(lib)
        prims = np.ascontiguousarray(prims, dtype=np.float64).reshape(-1, 16)
        mats = np.ascontiguousarray(mats, dtype=np.float64).reshape(-1, 8)
        texs = np.ascontiguousarray(texs, dtype=np.float64).reshape(-1, 8)
        h = C.c_void_p()
        rc = lib.rth_scene_from_records(_ptr(prims), prims.shape[0], _ptr(mats), mats.shape[0], _ptr(texs), texs.shape[0],
                                        image_paths.encode(), top_budget, C.byref(h))
        _check(lib, rc, host=True)
        return cls(h, lib)

    def close(self):
        if self._h:
            self._lib.rth_scene_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def desc(self) -> RtbSceneDesc:
        return self._lib.rth_scene_desc(self._h).contents

    def tree_info(self) -> dict:
        """rtb_tree_info of the scene's compiled device tree (host work, no GPU): sizes, stack bounds and the self-check counts."""
        if not self.desc.compiled:
            raise RuntimeError("the scene carries no compiled device tree (host_only / device_tree build)")
        info = (C.c_uint32 * 8)()
        cuda = self._lib._cuda
        cuda.rtb_tree_info.argtypes = [C.c_void_p, C.POINTER(C.c_uint32)]
        if cuda.rtb_tree_info(self.desc.compiled, info) != 0:
            raise RuntimeError("rtb_tree_info failed")
        keys = ("pairs", "pair_depth", "pair_prims", "pair_box_escapes", "quads", "quad_stack", "quad_prims", "quad_box_escapes")
        return dict(zip(keys, (int(v) for v in info)))

    def export(self):
        """(prims[n,16], mats[n,8], texs[n,8], image_paths) — the oracle's input."""
        pp, pm, pt = C.c_void_p(), C.c_void_p(), C.c_void_p()
        n_p, n_m, n_t = C.c_size_t(), C.c_size_t(), C.c_size_t()
        img = C.c_char_p()
        _check(self._lib, self._lib.rth_scene_export(self._h, C.byref(pp), C.byref(n_p), C.byref(pm), C.byref(n_m), C.byref(pt),
                                                     C.byref(n_t), C.byref(img)), host=True)

        def arr(ptr, n, w):
            if not n:
                return np.zeros((0, w))
            return np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_double)), shape=(n * w,)).reshape(n, w).copy()

        return arr(pp, n_p.value, 16), arr(pm, n_m.value, 8), arr(pt, n_t.value, 8), (img.value or b"").decode()

    def gid_to_ref(self) -> np.ndarray:
        p, n = C.c_void_p(), C.c_size_t()
        _check(self._lib, self._lib.rth_scene_gid_map(self._h, C.byref(p), C.byref(n)), host=True)
        if not n.value:
            return np.zeros(0, np.uint32)
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint32)), shape=(n.value,)).copy()

    def bvh_count(self) -> int:
        return self._lib.rth_scene_bvh_count(self._h)

    def bvh_dump(self, k: int):
        nn, npr = C.c_size_t(), C.c_size_t()
        _check(self._lib, self._lib.rth_scene_bvh_size(self._h, k, C.byref(nn), C.byref(npr)), host=True)
        nodes = np.zeros((nn.value, 9), np.float64)
        idx = np.zeros(npr.value, np.int32)
        _check(self._lib, self._lib.rth_scene_bvh_dump(self._h, k, _ptr(nodes), _ptr(idx)), host=True)
        return nodes, idx

    def nodes(self) -> np.ndarray:
        d = self.desc
        dt = np.dtype([("lo", np.float32, 3), ("a", np.uint32), ("hi", np.float32, 3), ("b", np.uint32)])
        return np.ctypeslib.as_array(C.cast(d.nodes, C.POINTER(C.c_uint8)), shape=(d.n_nodes * 32,)).view(dt).copy()

    def prim_refs(self) -> np.ndarray:
        d = self.desc
        return np.ctypeslib.as_array(C.cast(d.prim_refs, C.POINTER(C.c_uint32)), shape=(d.n_prim_refs,)).copy()


def camera_config(**kw) -> RthCameraConfig:
    """scene::CameraConfig with the reference's defaults (src/scene/camera.h:25-38)."""
    c = RthCameraConfig()
    c.aspect_ratio = kw.get("aspect_ratio", 16 / 9.0)
    c.image_width = kw.get("image_width", 400)
    c.samples_per_pixel = kw.get("samples_per_pixel", 50)
    c.max_depth = kw.get("max_depth", 10)
    c.vfov = kw.get("vfov", 90.0)
    c.lookfrom[:] = kw.get("lookfrom", (0, 0, 0))
    c.lookat[:] = kw.get("lookat", (0, 0, -1))
    c.vup[:] = kw.get("vup", (0, 1, 0))
    c.defocus_angle = kw.get("defocus_angle", 0.0)
    c.focus_dist = kw.get("focus_dist", 10.0)
    return c


def camera_preset(json_path, name: str, counters: bool = False) -> RthCameraConfig:
    lib = load_host()
    c = RthCameraConfig()
    _check(lib, lib.rth_camera_preset(str(json_path).encode(), name.encode(), C.byref(c)), host=True)
    return c


def encode_png(rgb8: np.ndarray, counters: bool = False) -> bytes:
    """rgb8: uint8 [H, W, 3] (what `resolve` returns) -> the bytes of a PNG file (rth_encode_png)."""
    lib = load_host()
    img = np.ascontiguousarray(rgb8, dtype=np.uint8)
    if img.ndim != 3 or img.shape[2] != 3:
        raise ValueError("encode_png: expected uint8 [H, W, 3]")
    h, w = img.shape[:2]
    n = C.c_size_t(0)
    _check(lib, lib.rth_encode_png(_ptr(img), w, h, None, 0, C.byref(n)), host=True)
    out = np.empty(n.value, dtype=np.uint8)
    _check(lib, lib.rth_encode_png(_ptr(img), w, h, _ptr(out), out.size, C.byref(n)), host=True)
    return out[: n.value].tobytes()


def camera_init(cfg: RthCameraConfig, depth_camera: bool = False, counters: bool = False) -> RtbCamera:
    lib = load_host()
    cam = RtbCamera()
    _check(lib, lib.rth_camera_init(C.byref(cfg), int(depth_camera), C.byref(cam)), host=True)
    return cam


class DeviceScene:
    """RtbScene*: the scene resident on one GPU."""

    def __init__(self, host_scene: HostScene, device: int = 0):
        self._lib = host_scene._lib
        _check(self._lib, self._lib.rtb_init(device))
        self._h = C.c_void_p()
        desc = host_scene.desc
        _check(self._lib, self._lib.rtb_scene_create(C.byref(desc), C.byref(self._h)))

    def close(self):
        if self._h:
            self._lib.rtb_scene_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def intersect(self, rays: np.ndarray) -> np.ndarray:
        """seam 2 with host buffers: rays RAY_DTYPE[n] -> HIT_DTYPE[n]."""
        rays = np.ascontiguousarray(rays, dtype=RAY_DTYPE)
        hits = np.zeros(rays.shape[0], HIT_DTYPE)
        _check(self._lib, self._lib.rtb_intersect_batch(self._h, _ptr(rays), rays.shape[0], _ptr(hits)))
        return hits

    def intersect_static(self, rays: np.ndarray) -> np.ndarray:
        """rtb_intersect_batch_static: seam 2 without any allocation on the device (at most STATIC_BATCH_MAX rays)."""
        rays = np.ascontiguousarray(rays, dtype=RAY_DTYPE)
        hits = np.zeros(rays.shape[0], HIT_DTYPE)
        self._lib.rtb_intersect_batch_static.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
        _check(self._lib, self._lib.rtb_intersect_batch_static(self._h, _ptr(rays), rays.shape[0], _ptr(hits)))
        return hits

    def intersect_device(self, d_rays_ptr: int, n: int, d_hits_ptr: int, stream: int = 0):
        _check(self._lib, self._lib.rtb_intersect_batch_device(self._h, d_rays_ptr, n, d_hits_ptr, stream))

    def render(self, cam: RtbCamera, spp: int, max_depth: int, mode: int = MODE_A, seed: int = 0x3360, spp_begin: int = 0,
               pool_size: int = 0, flags: int = 0):
        """seam 1 with a host buffer: returns (rgb_sum[h,w,3] float32, stats dict)."""
        prm = RtbRenderParams(spp_begin, spp_begin + spp, max_depth, mode, seed, pool_size, flags)
        out = np.zeros((cam.height, cam.width, 3), np.float32)
        st = RtbStats()
        _check(self._lib, self._lib.rtb_render(self._h, C.byref(cam), C.byref(prm), _ptr(out), C.byref(st)))
        return out, st.as_dict()

    def render_adaptive(self, cam: RtbCamera, max_spp: int, max_depth: int, rel_threshold: float = 0.05, min_samples: int = 16,
                        criterion: int = ADAPT_PIXEL_STATE, mode: int = MODE_B, seed: int = 0x3360, spp_begin: int = 0, flags: int = 0,
                        state=None):
        """seam 1 with the per-pixel adaptive stop; host buffers.  Returns (rgb_sum[h,w,3], rgb_sumsq[h,w,3], count[h,w], stats).
        `state` = (rgb_sum, rgb_sumsq, count) of an earlier call to continue (sets RENDER_KEEP_ACCUM)."""
        prm = RtbRenderParams(spp_begin, spp_begin + max_spp, max_depth, mode, seed, 0, flags | (RENDER_KEEP_ACCUM if state is not None else 0))
        ad = RtbAdaptive(rel_threshold, min_samples, criterion, 0)
        shape = (cam.height, cam.width)
        if state is not None:
            out, sq, cnt = (np.array(state[0], np.float32, order="C"), np.array(state[1], np.float32, order="C"), np.array(state[2], np.uint32, order="C"))
        else:
            out, sq, cnt = np.zeros(shape + (3,), np.float32), np.zeros(shape + (3,), np.float32), np.zeros(shape, np.uint32)
        st = RtbStats()
        _check(self._lib, self._lib.rtb_render_adaptive(self._h, C.byref(cam), C.byref(prm), C.byref(ad), _ptr(out), _ptr(sq), _ptr(cnt), C.byref(st)))
        return out, sq, cnt, st.as_dict()

    def render_adaptive_device(self, cam: RtbCamera, d_sum_ptr: int, d_sumsq_ptr: int, d_count_ptr: int, spp_begin: int, spp_end: int, max_depth: int,
                               rel_threshold: float = 0.05, min_samples: int = 16, criterion: int = ADAPT_PIXEL_STATE, mode: int = MODE_B,
                               seed: int = 0x3360, flags: int = 0):
        prm = RtbRenderParams(spp_begin, spp_end, max_depth, mode, seed, 0, flags)
        ad = RtbAdaptive(rel_threshold, min_samples, criterion, 0)
        st = RtbStats()
        _check(self._lib, self._lib.rtb_render_adaptive_device(self._h, C.byref(cam), C.byref(prm), C.byref(ad), d_sum_ptr, d_sumsq_ptr, d_count_ptr,
                                                               C.byref(st)))
        return st.as_dict()

    def render_device(self, cam: RtbCamera, d_accum_ptr: int, spp_begin: int, spp_end: int, max_depth: int, mode: int = MODE_A,
                      seed: int = 0x3360, pool_size: int = 0, flags: int = 0):
        prm = RtbRenderParams(spp_begin, spp_end, max_depth, mode, seed, pool_size, flags)
        st = RtbStats()
        _check(self._lib, self._lib.rtb_render_device(self._h, C.byref(cam), C.byref(prm), d_accum_ptr, C.byref(st)))
        return st.as_dict()


def render_multi(host_scene: "HostScene", cam: RtbCamera, spp: int, max_depth: int, n_devices: int, devices=None, mode: int = MODE_A,
                 seed: int = 0x3360, spp_begin: int = 0, flags: int = 0):
    """seam 1 on several GPUs from one host thread (rtb_render_multi): (rgb_sum[h,w,3], stats)."""
    lib = host_scene._lib
    prm = RtbRenderParams(spp_begin, spp_begin + spp, max_depth, mode, seed, 0, flags)
    out = np.zeros((cam.height, cam.width, 3), np.float32)
    st = RtbStats()
    dev = np.asarray(devices, np.int32) if devices is not None else None
    desc = host_scene.desc
    _check(lib, lib.rtb_render_multi(C.byref(desc), n_devices, _ptr(dev) if dev is not None else None, C.byref(cam), C.byref(prm), _ptr(out),
                                     C.byref(st)))
    return out, st.as_dict()


def pixel_state_fold(samples: np.ndarray, rel_threshold: float = 0.05, min_samples: int = 16, criterion: int = ADAPT_PIXEL_STATE,
                     counters: bool = False):
    """integrator::RecordSample + IsConverged over given sample sequences [n, k, 3]: (sum[n,3], sumsq[n,3], count[n])."""
    lib = load(counters)
    a = np.ascontiguousarray(samples, dtype=np.float32)
    n, k, _ = a.shape
    ad = RtbAdaptive(rel_threshold, min_samples, criterion, 0)
    s, q, c = np.zeros((n, 3), np.float32), np.zeros((n, 3), np.float32), np.zeros(n, np.uint32)
    _check(lib, lib.rtb_pixel_state_fold(_ptr(a), n, k, C.byref(ad), _ptr(s), _ptr(q), _ptr(c)))
    return s, q, c


def resolve_counts(rgb_sum: np.ndarray, count: np.ndarray, counters: bool = False) -> np.ndarray:
    """write_color of rgb_sum / count per pixel (wavefront.cc:229-241)."""
    lib = load(counters)
    a = np.ascontiguousarray(rgb_sum, dtype=np.float32)
    c = np.ascontiguousarray(count, dtype=np.uint32)
    out = np.zeros(a.shape, np.uint8)
    _check(lib, lib.rtb_resolve_counts(_ptr(a), _ptr(c), c.size, _ptr(out)))
    return out


def resolve(rgb_sum: np.ndarray, spp: int, counters: bool = False) -> np.ndarray:
    lib = load(counters)
    a = np.ascontiguousarray(rgb_sum, dtype=np.float32)
    out = np.zeros(a.shape, np.uint8)
    _check(lib, lib.rtb_resolve(_ptr(a), a.size // 3, spp, _ptr(out)))
    return out


def check_failures(counters: bool = False):
    """(is_checked_build, counts[8]) of the library's own bounds checker (rtb_debug_check_failures)."""
    lib = load(counters)
    c = np.zeros(8, np.uint64)
    rc = lib.rtb_debug_check_failures(_ptr(c))
    if rc < 0:
        _check(lib, rc)
    return rc == 1, c


def measure_peaks(counters: bool = False, l2_mib=(64,)) -> dict:
    """Roofline denominators measured on this GPU, now: L2 read bandwidth (ld.global.cg sweep of a 64 MiB buffer - inside the
    126 MB L2, far outside the L1s; other footprints in l2_gbs_by_mib), FP32 and FP64 FMA rates."""
    lib = load(counters)
    l2, f32, f64 = C.c_float(), C.c_float(), C.c_float()
    by = {}
    for mib in l2_mib:
        _check(lib, lib.rtb_measure_l2_gbs(int(mib) << 20, 40, C.byref(l2)))
        by[int(mib)] = l2.value
    _check(lib, lib.rtb_measure_fp32_tflops(20000, C.byref(f32)))
    _check(lib, lib.rtb_measure_fp64_tflops(20000, C.byref(f64)))
    return {"l2_gbs": by[int(l2_mib[0])], "l2_gbs_by_mib": by, "fp32_tflops": f32.value, "fp64_tflops": f64.value}


def write_ppm(path, rgb8: np.ndarray):
    """ASCII P3, the reference's output format (mega_kernel.h:48-51, color.h:32)."""
    h, w, _ = rgb8.shape
    with open(path, "w") as f:
        f.write(f"P3\n{w} {h}\n255\n")
        np.savetxt(f, rgb8.reshape(-1, 3), fmt="%d")
