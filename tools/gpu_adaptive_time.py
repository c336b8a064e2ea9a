"""Adaptive and fixed-spp render times of the bunny configuration, both estimators, several repetitions."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import bench, __graft_entry__ as g, numpy as np
m = g.load_package()
cfg = bench.CONFIGS["bunny"]
cam = m.camera_init(bench.camera_for(m, cfg))
ds = m.DeviceScene(m.HostScene.build(*cfg["scene"]))
for mode, nm in ((m.MODE_B, "B"), (m.MODE_A, "A")):
    ts = []
    for rep in range(8):
        s, q, c, st = ds.render_adaptive(cam, max_spp=cfg["spp"], max_depth=cfg["depth"], mode=mode)
        ts.append(st["ms_total"])
    print("adaptive mode", nm, "ms", np.round(ts, 2), "samples", st["samples"], "mean spp", c.mean())
    ts = []
    for rep in range(5):
        img, st = ds.render(cam, spp=cfg["spp"], max_depth=cfg["depth"], mode=mode)
        ts.append(st["ms_total"])
    print("fixed mode", nm, "ms", np.round(ts, 2))
