import sys, os
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import bench, __graft_entry__ as g
m = g.load_package()
name = sys.argv[1] if len(sys.argv) > 1 else "bunny"
cfg = bench.CONFIGS[name]
hs = m.HostScene.build(*cfg["scene"]); ds = m.DeviceScene(hs)
cam = m.camera_init(bench.camera_for(m, cfg))
ds.render(cam, spp=cfg["spp"], max_depth=cfg["depth"], flags=m.RENDER_SERIAL)
os.environ["RTB_TRACE_ITERS"] = "1"
ds.render(cam, spp=cfg["spp"], max_depth=cfg["depth"], flags=m.RENDER_SERIAL)
