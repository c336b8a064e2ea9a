"""C5 stress scene (4.0M triangles + 1M spheres, 3840x2160): throughput at reduced spp, both pipelines, counters."""
import sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import bench, __graft_entry__ as g
m = g.load_package()
cfg = bench.CONFIGS["stress"]
t = time.time(); hs = m.HostScene.build(*cfg["scene"]); print("host build %.1f s" % (time.time() - t), flush=True)
t = time.time(); ds = m.DeviceScene(hs); print("upload %.2f s" % (time.time() - t), flush=True)
cam = m.camera_init(bench.camera_for(m, cfg))
for spp in [int(a) for a in sys.argv[1:]] or [4, 16]:
    for label, flags in (("path", 0), ("wavefront", m.RENDER_WAVEFRONT)):
        ts = []
        for rep in range(3):
            img, st = ds.render(cam, spp=spp, max_depth=cfg["depth"], flags=flags)
            ts.append(st["ms_total"])
        tt = min(ts[1:])
        print("spp %4d %-9s %9.1f ms %8.1f Msamples/s %8.1f Mrays/s rays/sample %.2f mean %.4f" % (
            spp, label, tt, st["samples"] / tt / 1e3, st["rays"] / tt / 1e3, st["rays"] / st["samples"], float(img.mean()) / spp), flush=True)
m.write_ppm  # keep a picture: 1/4 scale box filter
sm = (img / spp).reshape(cam.height // 4, 4, cam.width // 4, 4, 3).mean(axis=(1, 3))
m.write_ppm("gpurun_out/stress_quarter.ppm", m.resolve((sm * 1).astype(np.float32), 1))
hs_c = m.HostScene.build(*cfg["scene"], counters=True); ds_c = m.DeviceScene(hs_c)
_, sc = ds_c.render(cam, spp=1, max_depth=cfg["depth"])
print("per ray: nodes %.2f sph %.2f tri %.2f" % (sc["node_visits"] / sc["rays"], sc["sphere_tests"] / sc["rays"], sc["triangle_tests"] / sc["rays"]))
