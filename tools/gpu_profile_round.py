"""Short program for ncu (session 17): scene creation with the device tree build, one fixed-spp render and one
adaptive render of the bunny configuration at reduced spp."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import bench, __graft_entry__ as g
m = g.load_package()
cfg = bench.CONFIGS["bunny"]
cam = m.camera_init(bench.camera_for(m, cfg))
for dev in (False, True):
    ds = m.DeviceScene(m.HostScene.build(*cfg["scene"], device_tree=dev))
    for rep in range(2):
        img, st = ds.render(cam, spp=cfg["spp"], max_depth=cfg["depth"])
    print("tree", "device" if dev else "host", "fixed", st["ms_total"], st["rays"])
    s, q, c, st = ds.render_adaptive(cam, max_spp=cfg["spp"], max_depth=cfg["depth"], mode=m.MODE_B)
    print("adaptive", st["ms_total"], st["samples"], c.mean())
    ds.close()
