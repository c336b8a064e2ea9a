"""Small workload for compute-sanitizer (memcheck / racecheck / initcheck): every kernel of the library once, on the mixed test
scene (spheres + triangles + rects, nested BVHs): a fixed-spp render with the path kernel, the pool kernel and the wavefront
kernels (generate / extend / shade + drain), an adaptive render (path kernel + replay), rtb_intersect_batch (+ exact_t),
the device tree build, resolve.  Sizes are tiny: the tools slow kernels down 10-100x.
  compute-sanitizer --tool memcheck python tools/gpu_sanitize.py"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))
import helpers  # noqa: E402
import __graft_entry__ as g  # noqa: E402

m = g.load_package()
hs = m.HostScene.from_records(*helpers.mixed_scene())
ds = m.DeviceScene(hs)
cam = m.camera_init(m.camera_config(image_width=64, vfov=60.0, lookfrom=(1, 2.5, 7), lookat=(0, 1, 0)))
ref, st = ds.render(cam, spp=4, max_depth=8)
print("path kernel", st["rays"], float(ref.sum()))
for label, flags, kw in (("pool kernel", m.RENDER_POOL_PATHS, {}), ("wavefront", m.RENDER_WAVEFRONT, {"pool_size": 1 << 11}),
                         ("wavefront serial", m.RENDER_SERIAL, {"pool_size": 1 << 11}), ("smem top", m.RENDER_SMEM_TOP, {})):
    img, s2 = ds.render(cam, spp=4, max_depth=8, flags=flags, **kw)
    assert s2["rays"] == st["rays"] and np.abs(img - ref).max() <= 1e-4 * max(1.0, ref.max()), label
    print(label, s2["rays"], s2["launches"])
for mode in (m.MODE_A, m.MODE_B):
    s, q, c, sa = ds.render_adaptive(cam, max_spp=24, max_depth=8, mode=mode)
    print("adaptive", mode, sa["samples"], float(c.mean()))
rays = helpers.ray_batch(m, 4096, 5)
h = ds.intersect(rays)
print("intersect_batch", int((h["prim_id"] >= 0).sum()))
dev = m.DeviceScene(m.HostScene.build("stress", (24, 23, 12, 20, 2), device_tree=True))
h2 = dev.intersect(helpers.ray_batch(m, 2048, 6, eye=(0, 20, 30), spread=5.0))
img, s3 = dev.render(cam, spp=2, max_depth=6)
print("device tree", int((h2["prim_id"] >= 0).sum()), s3["rays"])
rgb8 = m.resolve(ref, 4)
print("resolve", int(rgb8.sum()))
print("sanitize workload done")
