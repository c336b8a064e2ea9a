"""What the cost-ordered schedule is worth (bunny C3 and RTIOW C1): plain order; measuring schedule (3 overlapping launches);
and - the ordering alone - one launch in the order a previous render measured (RTB_COST_ORDER_REUSE)."""
import os, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import bench, __graft_entry__ as g
m = g.load_package()
for name in sys.argv[1:] or ["bunny", "rtiow"]:
    cfg = bench.CONFIGS[name]
    ds = m.DeviceScene(m.HostScene.build(*cfg["scene"]))
    cam = m.camera_init(bench.camera_for(m, cfg))
    def run(label, spp, env):
        for k in ("RTB_COST_ORDER", "RTB_COST_ORDER_REUSE"):
            os.environ.pop(k, None)
        os.environ.update(env)
        ts = []
        for rep in range(7):
            img, st = ds.render(cam, spp=spp, max_depth=cfg["depth"])
            ts.append(st["ms_total"])
        print(f"{name:8s} {label:34s} spp {spp:4d}: median {np.median(ts[2:]):7.3f} ms  min {min(ts):7.3f}  launches {st['launches']}", flush=True)
    for spp in (cfg["spp"], max(8, cfg["spp"] // 8)):
        run("plain", spp, {})
        run("measuring schedule (3 launches)", spp, {"RTB_COST_ORDER": "1"})
        run("one launch, order of previous render", spp, {"RTB_COST_ORDER": "1", "RTB_COST_ORDER_REUSE": "1"})
    for depth_spp in (1,):
        run("plain 1 spp (the launch floor)", 1, {})
    ds.close()
