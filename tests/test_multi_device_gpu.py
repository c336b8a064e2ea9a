"""rtb_render_multi: one host thread, several GPUs of one box (worker thread per device, spp slices of the global
sample indices, peer-access reduce on the first device).  Needs >= 2 visible devices: skipped on a single-GPU box
(run with `gpurun --gpus 2 -- python -m pytest tests/test_multi_device_gpu.py`)."""
import subprocess
from pathlib import Path

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parents[1]


@pytest.fixture(scope="module")
def n_dev(rt):
    n = rt.load().rtb_device_count()
    if n < 2:
        pytest.skip("needs at least two CUDA devices")
    return n


def test_multi_equals_single_device_render(rt, n_dev):
    hs = rt.HostScene.build("rtiow", (11, 1))
    cam = rt.camera_init(rt.camera_config(aspect_ratio=1.777, image_width=320, vfov=45.0, lookfrom=(0, 2, 5), lookat=(0, 0, 0)))
    ds = rt.DeviceScene(hs)
    one, s1 = ds.render(cam, spp=37, max_depth=10)
    for n in sorted({2, min(n_dev, 3), n_dev}):
        multi, sm = rt.render_multi(hs, cam, spp=37, max_depth=10, n_devices=n)
        assert sm["samples"] == s1["samples"] and sm["rays"] == s1["rays"], f"{n} devices trace other paths"
        assert np.abs(multi - one).max() <= 3e-5 * one.max()
    # an explicit device list, reversed: the root of the reduce is the last device
    multi, sm = rt.render_multi(hs, cam, spp=37, max_depth=10, n_devices=2, devices=[n_dev - 1, 0])
    assert sm["rays"] == s1["rays"] and np.abs(multi - one).max() <= 3e-5 * one.max()


def test_multi_argument_errors(rt, n_dev):
    hs = rt.HostScene.build("cornell")
    cam = rt.camera_init(rt.camera_config(image_width=32))
    with pytest.raises(rt.RtbError):
        rt.render_multi(hs, cam, spp=4, max_depth=4, n_devices=2, devices=[0, 0])
    with pytest.raises(rt.RtbError):
        rt.render_multi(hs, cam, spp=4, max_depth=4, n_devices=2, devices=[0, n_dev])
    with pytest.raises(rt.RtbError):
        rt.render_multi(hs, cam, spp=4, max_depth=4, n_devices=0)


def test_cli_gpus_flag(rt, n_dev):
    exe = ROOT / "3360-ray-tracer_b200" / "rt_b200"
    base = [str(exe), "cornell", "--scene", "cornell", "--width", "60", "--spp", "64", "--depth", "8", "--assets", str(ROOT / "assets"),
            "--cameras", str(ROOT / "assets" / "cameras.json")]
    a = subprocess.run(base, capture_output=True, cwd=ROOT, timeout=300)
    b = subprocess.run(base + ["--gpus", "2"], capture_output=True, cwd=ROOT, timeout=300)
    assert a.returncode == 0 and b.returncode == 0, b.stderr.decode()[-400:]
    pa = np.array(a.stdout.split()[4:], dtype=np.int32)
    pb = np.array(b.stdout.split()[4:], dtype=np.int32)
    assert pa.shape == pb.shape and np.abs(pa - pb).max() <= 1 and (pa != pb).mean() < 0.01
