"""Render the golden-fixture cameras on the GPU at high spp and dump float images for offline comparison."""
import sys
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import numpy as np
import __graft_entry__ as g
m = g.load_package()
from test_host_scene import SCENES
from test_render_gpu import golden_camera
out = {}
for name in (sys.argv[1:] or list(SCENES)):
    gd = np.load(f'tests/golden/image_{name}.npz')
    cfg, cam = golden_camera(m, gd)
    hs = SCENES[name](m); ds = m.DeviceScene(hs)
    img, st = ds.render(cam, spp=1024, max_depth=cfg.max_depth, seed=77)
    out[name] = img / 1024
    print(name, st)
np.savez_compressed('gpurun_out/gpu_images.npz', **out)
