"""Per-launch constant of the path kernel: fixed-spp renders of the bunny configuration at small spp."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import bench, __graft_entry__ as g
m = g.load_package()
cfg = bench.CONFIGS["bunny"]
cam = m.camera_init(bench.camera_for(m, cfg))
ds = m.DeviceScene(m.HostScene.build(*cfg["scene"]))
for mode, mn in ((m.MODE_A, "A"), (m.MODE_B, "B")):
    for spp in (1, 2, 4, 8, 16, 32, 100):
        ts = [ds.render(cam, spp=spp, max_depth=cfg["depth"], mode=mode)[1]["ms_total"] for _ in range(5)]
        print(f"mode {mn} spp {spp:3d}: {np.median(ts[1:]):7.3f} ms  ({360000 * spp / np.median(ts[1:]) / 1e3:7.1f} Msamples/s)", flush=True)
for chunk in (8, 16):
    import os
    os.environ["RTB_ADAPT_CHUNK"] = str(chunk)
    for thr, mnq in ((0.05, 16), (0.0, 100)):
        ts = []
        for _ in range(4):
            s, q, c, st = ds.render_adaptive(cam, max_spp=100, max_depth=cfg["depth"], mode=m.MODE_B, rel_threshold=thr, min_samples=mnq)
            ts.append(st["ms_total"])
        print(f"adaptive chunk {chunk} thr {thr} min {mnq}: {np.median(ts[1:]):7.3f} ms, started {st['samples']}, recorded {int(c.sum())}, launches {st['launches']}", flush=True)
