"""Short program for ncu: two full renders of one BASELINE config (the first warms pools and caches).
  python tools/gpu_profile_step.py [config] [spp] [mode]
"""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import bench  # noqa: E402
import __graft_entry__ as g  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "bunny"
m = g.load_package()
cfg = bench.CONFIGS[name]
spp = int(sys.argv[2]) if len(sys.argv) > 2 else cfg["spp"]
mode = m.MODE_B if len(sys.argv) > 3 and sys.argv[3] == "B" else m.MODE_A
hs = m.HostScene.build(*cfg["scene"])
ds = m.DeviceScene(hs)
cam = m.camera_init(bench.camera_for(m, cfg, None, spp))
for rep in range(2):
    img, st = ds.render(cam, spp=spp, max_depth=cfg["depth"], mode=mode)
    print(name, "rep", rep, {k: st[k] for k in ("samples", "rays", "iterations", "launches", "ms_total")},
          "Msamples/s %.1f" % (st["samples"] / st["ms_total"] / 1e3), flush=True)
