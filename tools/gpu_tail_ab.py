"""Tail hand-over A/B (bunny, rtiow, cornell): full spp and 1/8 of it (one of 8 GPUs in the spp split), 1 spp (the launch floor)."""
import os, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import bench, __graft_entry__ as g
m = g.load_package()
for name in sys.argv[1:] or ["bunny", "rtiow", "cornell", "earth"]:
    cfg = bench.CONFIGS[name]
    ds = m.DeviceScene(m.HostScene.build(*cfg["scene"]))
    cam = m.camera_init(bench.camera_for(m, cfg))
    ref = None
    for label, env in (("hand-over", {}), ("no hand-over", {"RTB_NO_TAIL_HANDOVER": "1"})):
        os.environ.pop("RTB_NO_TAIL_HANDOVER", None)
        os.environ.update(env)
        for spp in (cfg["spp"], max(8, cfg["spp"] // 8), 1):
            ts = []
            for rep in range(7):
                img, st = ds.render(cam, spp=spp, max_depth=cfg["depth"])
                ts.append(st["ms_total"])
            key = (spp,)
            print(f"{os.environ.get('RTB_LIB', 'librtb200.so'):20s} {name:8s} {label:13s} spp {spp:4d}: median {np.median(ts[2:]):7.3f} ms  min {min(ts):7.3f}  launches {st['launches']} rays {st['rays']} sum {img.astype(np.float64).sum():.6e}", flush=True)
    ds.close()
