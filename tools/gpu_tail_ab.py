"""Tail A/B: every config at its spp, at 1/8 of it (one of 8 GPUs in the spp split) and at 1 spp (the launch floor); prints a checksum
of the image so that two libraries (RTB_LIB) can be compared."""
import os, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import bench, __graft_entry__ as g
m = g.load_package()
for name in sys.argv[1:] or ["bunny", "rtiow", "cornell", "earth"]:
    cfg = bench.CONFIGS[name]
    ds = m.DeviceScene(m.HostScene.build(*cfg["scene"], device_tree=name == "stress"))
    cam = m.camera_init(bench.camera_for(m, cfg))
    full = 8 if name == "stress" else cfg["spp"]
    for spp in (full, max(1, full // 8), 1):
        ts = []
        for rep in range(7):
            img, st = ds.render(cam, spp=spp, max_depth=cfg["depth"])
            ts.append(st["ms_total"])
        print(f"{os.environ.get('RTB_LIB', 'librtb200.so'):24s} {name:8s} spp {spp:4d}: median {np.median(ts[2:]):8.3f} ms  min {min(ts):8.3f}  rays {st['rays']} sum {img.astype(np.float64).sum():.9e}", flush=True)
    ds.close()
