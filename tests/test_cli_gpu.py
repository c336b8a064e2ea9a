"""rt_b200, the reference's main() (src/main.cc:158-197) on the device path: PPM on stdout in the reference's
format, the binary variant, the progressive accumulator, the adaptive renderer."""
import subprocess
from pathlib import Path

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parents[1]
EXE = ROOT / "3360-ray-tracer_b200" / "rt_b200"


def run(*args):
    r = subprocess.run([str(EXE), *args, "--assets", str(ROOT / "assets"), "--cameras", str(ROOT / "assets" / "cameras.json")],
                       capture_output=True, cwd=ROOT, timeout=300)
    assert r.returncode == 0, r.stderr.decode()[-500:]
    return r.stdout, r.stderr.decode()


def parse_p3(data):
    tok = data.split()
    assert tok[0] == b"P3" and tok[3] == b"255"
    w, h = int(tok[1]), int(tok[2])
    px = np.array(tok[4:], dtype=np.int32)
    assert px.size == w * h * 3 and px.min() >= 0 and px.max() <= 255
    return px.reshape(h, w, 3)


def parse_p6(data):
    head, rest = data.split(b"\n255\n", 1)
    magic, dims = head.split(b"\n")
    assert magic == b"P6"
    w, h = (int(x) for x in dims.split())
    assert len(rest) == w * h * 3
    return np.frombuffer(rest, np.uint8).reshape(h, w, 3).astype(np.int32)


def test_p3_output_and_runtime_line():
    out, err = run("cornell", "--scene", "cornell", "--width", "60", "--spp", "64", "--depth", "8")
    img = parse_p3(out)
    assert img.shape == (40, 60, 3) and img.mean() > 5  # 60 / aspect 1.5
    assert "Runtime:" in err and "Done." in err
    # one line per pixel "r g b" like core::write_color (color.h:32)
    lines = out.decode().splitlines()
    assert lines[0] == "P3" and lines[1] == "60 40" and lines[2] == "255" and len(lines) == 3 + 60 * 40


def test_p6_carries_the_same_bytes():
    args = ("cornell", "--scene", "cornell", "--width", "60", "--spp", "64", "--depth", "8")
    a = parse_p3(run(*args)[0])
    b = parse_p6(run(*args, "--p6")[0])
    # same Philox samples; only the float order of the atomic adds differs between two runs
    assert np.abs(a - b).max() <= 1 and (a != b).mean() < 0.01


def test_png_carries_the_same_bytes():
    from test_png import decode_png
    args = ("cornell", "--scene", "cornell", "--width", "60", "--spp", "64", "--depth", "8")
    a = parse_p3(run(*args)[0])
    b = decode_png(run(*args, "--png")[0]).astype(np.int32)
    assert b.shape == (40, 60, 3)
    assert np.abs(a - b).max() <= 1 and (a != b).mean() < 0.01


def test_dump_and_resume(tmp_path):
    base = ("cornell", "--scene", "cornell", "--width", "60", "--depth", "8")
    acc = tmp_path / "acc.bin"
    run(*base, "--spp", "24", "--dump", str(acc))
    resumed = parse_p3(run(*base, "--spp", "40", "--resume", str(acc))[0])
    full = parse_p3(run(*base, "--spp", "64")[0])
    assert np.abs(resumed - full).max() <= 1 and (resumed != full).mean() < 0.01


def test_adaptive_flag_and_mode_b():
    out, err = run("cornell", "--scene", "cornell", "--width", "60", "--spp", "64", "--depth", "8", "--mode", "B", "--adaptive", "0.05", "16")
    img = parse_p3(out)
    fixed = parse_p3(run("cornell", "--scene", "cornell", "--width", "60", "--spp", "64", "--depth", "8", "--mode", "B")[0])
    assert abs(img.mean() - fixed.mean()) < 6.0
    # adaptive renders start fewer camera samples than pixels x spp
    done = int(err.split("Total samples:")[1].split("Per")[0])
    assert 16 * 60 * 40 <= done <= 64 * 60 * 40
