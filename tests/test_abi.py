"""The C-ABI library loads and exports every symbol include/*.h declares (no compute without a GPU)."""
import ctypes as C
import re
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]


def declared(header):
    text = (ROOT / "include" / header).read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(rt[bh]_[a-z0-9_]+)\s*\(", text)))


def test_every_declared_symbol_is_exported(rt):
    lib = rt.load()
    names = declared("rtb.h") + declared("rth.h")
    assert len(names) >= 28
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/ but not exported by librtb200.so / librth.so"
    assert set(rt.binding.RTB_SYMBOLS) == set(declared("rtb.h"))
    assert set(rt.binding.RTH_SYMBOLS) == set(declared("rth.h"))


def test_the_two_libraries_are_separate(rt):
    """librth.so (host scene API) carries no rtb_* entry and needs neither CUDA nor librtb200.so; librtb200.so carries no
    rth_* entry.  `bench.py --impl reference` builds its scene through librth.so alone."""
    import subprocess
    host, cuda = rt.binding.load_host(), rt.load()._cuda
    for n in declared("rtb.h"):
        assert not hasattr(host, n), f"librth.so exports {n}"
    for n in declared("rth.h"):
        assert not hasattr(cuda, n), f"librtb200.so exports {n}"
    needed = subprocess.run(["objdump", "-p", str(rt.binding.host_lib_path())], capture_output=True, text=True).stdout
    assert "cudart" not in needed and "rtb200" not in needed
    code = ("import sys; sys.path.insert(0, %r); import __graft_entry__ as g; rt = g.load_package(); "
            "hs = rt.HostScene.build('cornell', host_only=True); assert hs.desc.n_rects == 6 and not hs.desc.compiled; "
            "import os; m = open('/proc/%%d/maps' %% os.getpid()).read(); assert 'librtb200' not in m and 'libcudart' not in m") % str(ROOT)
    subprocess.run([__import__("sys").executable, "-c", code], check=True, cwd=ROOT)


def test_counters_build_exports_the_same_abi(rt):
    lib = rt.load(counters=True)
    for n in declared("rtb.h"):
        assert hasattr(lib, n)


def test_abi_version_and_struct_sizes(rt):
    lib = rt.load()
    assert lib.rtb_abi_version() == 3
    assert C.sizeof(rt.binding.RtbNode) == 32
    assert rt.RAY_DTYPE.itemsize == 32 and rt.HIT_DTYPE.itemsize == 16
    assert C.sizeof(rt.binding.RtbCamera) == 88
    assert C.sizeof(rt.binding.RtbRenderParams) == 32


def test_no_cpu_fallback_without_device(rt):
    lib = rt.load()
    if lib.rtb_device_count() > 0:
        pytest.skip("a GPU is present")
    assert lib.rtb_init(0) == rt.RTB_ENODEVICE
    hs = rt.HostScene.build("cornell")
    with pytest.raises(rt.RtbError) as e:
        rt.DeviceScene(hs)
    assert e.value.code == rt.RTB_ENODEVICE and b"no CPU path" in lib.rtb_last_error()


def test_scene_create_rejects_malformed_descriptions(rt):
    lib = rt.load()
    hs = rt.HostScene.build("cornell")
    d = rt.binding.RtbSceneDesc.from_buffer_copy(hs.desc)
    d.abi_version = 99
    h = C.c_void_p()
    assert lib.rtb_scene_create(C.byref(d), C.byref(h)) == rt.RTB_EINVAL
    d = rt.binding.RtbSceneDesc.from_buffer_copy(hs.desc)
    d.n_nodes = 3
    assert lib.rtb_scene_create(C.byref(d), C.byref(h)) == rt.RTB_EINVAL
    assert lib.rtb_scene_create(None, C.byref(h)) == rt.RTB_EINVAL


def test_missing_library_fails_loudly(rt, monkeypatch, tmp_path):
    monkeypatch.setattr(rt.binding, "PKG_DIR", tmp_path)
    monkeypatch.setattr(rt.binding, "_LIBS", {})
    monkeypatch.setattr(rt.binding, "_HOST", [])
    with pytest.raises(ImportError):
        rt.binding.load()


def test_product_never_imports_the_oracle():
    pkg = ROOT / "3360-ray-tracer_b200"
    for p in list(pkg.rglob("*.py")) + list(pkg.rglob("*.h")) + list(pkg.rglob("*.cc")) + list(pkg.rglob("*.cu")) + list(pkg.rglob("*.cuh")):
        text = p.read_text(errors="ignore")
        assert "oracle" not in text.replace("for the oracle", "").replace("the oracle", "").replace("oracle export", "").replace("oracle's", "").lower() \
            or "import oracle" not in text and "from oracle" not in text and "liboracle" not in text, p


def test_sample_generator_division_is_exact(rt):
    """FastDiv (csrc/rtb_kernels.cuh): the multiply-shift division of the sample generator against Python's, for the divisors the
    configurations use (widths, pixel counts, samples per launch), the edges of the 32-bit range and a few thousand random pairs."""
    import numpy as np
    lib = rt.load()
    rng = np.random.default_rng(7)
    divisors = [1, 2, 3, 5, 7, 31, 32, 33, 44, 100, 200, 256, 450, 533, 562, 675, 800, 1000, 1024, 1200, 2160, 3840, 360000, 426400, 562000, 810000,
                8294400, (1 << 30) - 1, 1 << 30, (1 << 31) - 1, 1 << 31, (1 << 31) + 1, 0xfffffffe, 0xffffffff] + [int(v) for v in rng.integers(1, 1 << 32, 300)]
    for d in divisors:
        ns = [0, 1, d - 1, d, d + 1, 2 * d - 1, 2 * d, 0x7fffffff, 0x80000000, 0xfffffffe, 0xffffffff] + [int(v) for v in rng.integers(0, 1 << 32, 40)]
        for n in ns:
            n &= 0xffffffff
            assert lib.rtb_debug_fast_div(n, d) == n // d, (n, d)
    assert lib.rtb_debug_fast_div(5, 0) == 0
