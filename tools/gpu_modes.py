"""Wavefront (with / without drain) vs fused path kernel: time and image agreement on one config."""
import sys, os
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import bench, __graft_entry__ as g
m = g.load_package()
name = sys.argv[1] if len(sys.argv) > 1 else "bunny"
cfg = bench.CONFIGS[name]
spp = int(sys.argv[2]) if len(sys.argv) > 2 else cfg["spp"]
hs = m.HostScene.build(*cfg["scene"]); ds = m.DeviceScene(hs)
cam = m.camera_init(bench.camera_for(m, cfg))
def run(label, flags, env=None):
    for k, v in (env or {}).items():
        os.environ[k] = v
    ts = []
    for rep in range(5):
        img, st = ds.render(cam, spp=spp, max_depth=cfg["depth"], flags=flags)
        ts.append(st["ms_total"])
    for k in (env or {}):
        del os.environ[k]
    print("%-34s %.2f ms (min %.2f)  %.0f Msamples/s  rays %d launches %d drain_paths %d" % (
        label, np.median(ts[1:]), min(ts[1:]), st["samples"] / np.median(ts[1:]) / 1e3, st["rays"], st["launches"], st["drain_paths"]), flush=True)
    return img
ref = run("wavefront, no drain", m.RENDER_WAVEFRONT | m.RENDER_NO_DRAIN)
for dm in (65536, 262144, 524288, 1048576, 2097152):
    img = run("wavefront, drain <= %d" % dm, m.RENDER_WAVEFRONT, {"RTB_DRAIN_MAX": str(dm)})
    print("    max rel diff vs no-drain: %.2e" % (np.abs(img - ref).max() / ref.max()))
img = run("fused path kernel", 0)
print("    max rel diff vs wavefront: %.2e" % (np.abs(img - ref).max() / ref.max()))
