import sys, time, subprocess, threading
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import bench, __graft_entry__ as g
m = g.load_package()
cfg = bench.CONFIGS["stress"]
hs = m.HostScene.build(*cfg["scene"]); ds = m.DeviceScene(hs)
cam = m.camera_init(bench.camera_for(m, cfg))
clk = bench.ClockSampler(0); clk.start(); time.sleep(0.5)
for spp in (1, 2, 4, 8, 16, 4, 4, 4, 4, 32):
    t0 = time.perf_counter()
    img, st = ds.render(cam, spp=spp, max_depth=cfg["depth"])
    print("spp %3d %8.1f ms  %.1f Msamples/s  wall %.1f ms" % (spp, st["ms_total"], st["samples"] / st["ms_total"] / 1e3, 1e3 * (time.perf_counter() - t0)), flush=True)
time.sleep(0.2)
rows = clk.rows; clk.stop()
print("clock samples (sm MHz, W):", [(r[1][1], r[1][3]) for r in rows][::3])
