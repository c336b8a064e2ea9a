"""The torch.distributed / NCCL sharding of multigpu.py on real GPUs (one process per GPU under torchrun): the reduced
accumulator of a fixed-spp render equals the single-GPU render; an adaptive render reduces sums and counts.  Needs
>= 2 devices (skipped on the single-GPU box; run under `gpurun --gpus 2`)."""
import socket
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parents[1]


def test_nccl_sharded_render_equals_single_gpu(rt, tmp_path):
    if rt.load().rtb_device_count() < 2:
        pytest.skip("needs at least two CUDA devices")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    out = tmp_path / "nccl.npz"
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                        "--master-port", str(port), str(ROOT / "tests" / "nccl_worker.py"), str(out)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-800:]
    g = np.load(out)
    h, w = (int(x) for x in g["shape"])
    hs = rt.HostScene.build("rtiow", (11, 1))
    ds = rt.DeviceScene(hs)
    cam = rt.camera_init(rt.camera_config(aspect_ratio=1.777, image_width=192, vfov=45.0, lookfrom=(0, 2, 5), lookat=(0, 0, 0)))
    assert (cam.height, cam.width) == (h, w)
    one, st = ds.render(cam, spp=int(g["spp"]), max_depth=10)
    assert st["rays"] == int(g["rays"])
    assert np.abs(g["fixed"].reshape(h, w, 3) - one).max() <= 3e-5 * one.max()
    # adaptive, 2 ranks x (16 .. 64) samples per pixel: counts within bounds, image statistically the fixed render's
    cnt = g["a_cnt"].reshape(h, w)
    assert cnt.min() >= 32 and cnt.max() <= 128 and (cnt < 128).mean() > 0.05
    mean_a = g["a_sum"].reshape(h, w, 3) / cnt[..., None]
    mean_f = one / float(g["spp"])
    assert abs(mean_a.mean() - mean_f.mean()) <= 0.02 * mean_f.mean()
