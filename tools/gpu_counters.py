import sys, os
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
os.environ["RTB_DEBUG_COUNTERS"] = "1"
import bench, __graft_entry__ as g
m = g.load_package()
name = sys.argv[1] if len(sys.argv) > 1 else "bunny"
cfg = bench.CONFIGS[name]
hs = m.HostScene.build(*cfg["scene"], counters=True); ds = m.DeviceScene(hs)
cam = m.camera_init(bench.camera_for(m, cfg), counters=True)
img, st = ds.render(cam, spp=min(cfg["spp"], 8), max_depth=cfg["depth"])
print({k: st[k] for k in ("rays", "node_visits", "sphere_tests", "triangle_tests", "rect_tests")})
