"""All single-GPU BASELINE configs, wavefront vs fused: ms, Msamples/s, Mrays/s, image agreement."""
import sys, os
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import bench, __graft_entry__ as g
m = g.load_package()
names = sys.argv[1:] or ["rtiow", "cornell", "bunny", "earth"]
for name in names:
    cfg = bench.CONFIGS[name]
    hs = m.HostScene.build(*cfg["scene"]); ds = m.DeviceScene(hs); cam = m.camera_init(bench.camera_for(m, cfg))
    imgs = {}
    for label, flags in (("wavefront", m.RENDER_WAVEFRONT), ("fused", 0)):
        for mode, mname in ((m.MODE_A, "A"), (m.MODE_B, "B")):
            ts = []
            for rep in range(4):
                img, st = ds.render(cam, spp=cfg["spp"], max_depth=cfg["depth"], flags=flags, mode=mode)
                ts.append(st["ms_total"])
            t = float(np.median(ts[1:]))
            imgs[(label, mname)] = img
            print("%-8s %-9s mode %s %8.2f ms %8.1f Msamples/s %8.1f Mrays/s rays/sample %.2f" % (
                name, label, mname, t, st["samples"] / t / 1e3, st["rays"] / t / 1e3, st["rays"] / st["samples"]), flush=True)
    for mname in "AB":
        a, b = imgs[("wavefront", mname)], imgs[("fused", mname)]
        print("    mode %s fused vs wavefront max rel diff %.2e" % (mname, np.abs(a - b).max() / a.max()))
    ds.close()
