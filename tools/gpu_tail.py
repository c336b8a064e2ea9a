"""How the per-launch constant of the path kernel depends on the path depth and on the grid size (bunny, mode A)."""
import sys, os
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import bench, __graft_entry__ as g
m = g.load_package()
cfg = bench.CONFIGS["bunny"]
cam = m.camera_init(bench.camera_for(m, cfg))
ds = m.DeviceScene(m.HostScene.build(*cfg["scene"]))
for depth in (1, 2, 3, 5, 10, 20):
    row = []
    for spp in (1, 4, 16):
        ts = [ds.render(cam, spp=spp, max_depth=depth)[1] for _ in range(5)]
        row.append("spp %2d: %6.3f ms (%.2f rays/sample)" % (spp, np.median([t["ms_total"] for t in ts[1:]]), ts[-1]["rays"] / ts[-1]["samples"]))
    print("depth %2d  " % depth + "   ".join(row), flush=True)
