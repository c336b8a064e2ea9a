"""Short program for ncu: the stress configuration (device-built tree) at 2 spp, two renders."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import bench, __graft_entry__ as g
m = g.load_package()
cfg = bench.CONFIGS["stress"]
cam = m.camera_init(bench.camera_for(m, cfg))
ds = m.DeviceScene(m.HostScene.build(*cfg["scene"], device_tree=True))
for rep in range(2):
    img, st = ds.render(cam, spp=2, max_depth=cfg["depth"])
    print("stress 2 spp", st["ms_total"], st["rays"], flush=True)
