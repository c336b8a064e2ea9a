"""Host-built (SAH + device-cost leaf split) vs device-built (LBVH) tree: scene creation time and render time."""
import sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import bench, __graft_entry__ as g
m = g.load_package()
for name in sys.argv[1:] or ["bunny", "rtiow"]:
    cfg = bench.CONFIGS[name]
    spp = 4 if name == "stress" else cfg["spp"]
    cam = m.camera_init(bench.camera_for(m, cfg))
    for label, dev in (("host SAH tree ", False), ("device LBVH   ", True)):
        t0 = time.perf_counter()
        hs = m.HostScene.build(*cfg["scene"], device_tree=dev)
        t1 = time.perf_counter()
        ds = m.DeviceScene(hs)
        t2 = time.perf_counter()
        ds.close()
        t2b = time.perf_counter()
        ds = m.DeviceScene(hs)
        t3 = time.perf_counter()
        ts = []
        for rep in range(4):
            img, st = ds.render(cam, spp=spp, max_depth=cfg["depth"])
            ts.append(st["ms_total"])
        print(f"{name} {label}: host scene {t1-t0:.3f} s, rtb_scene_create {1e3*(t2-t1):.1f} ms (again {1e3*(t3-t2b):.1f} ms), render {np.median(ts[1:]):.2f} ms at {spp} spp, "
              f"{st['rays']/st['samples']:.3f} rays/sample, checksum {img.sum():.1f}", flush=True)
        ds.close()
