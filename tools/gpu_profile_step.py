"""Short program for ncu: two full renders of one BASELINE config (the first warms pools and caches).
  python tools/gpu_profile_step.py [config] [spp] [mode A|B] [flags: smemtop|pool|wavefront|adaptive|batch ...]
stress uses the device-built tree (what bench.py --config stress renders with).  `batch`: one rtb_intersect_batch of 2 M
camera rays instead of the renders; `adaptive`: one adaptive render (estimator B, 0.05 / 16)."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
sys.path.insert(0, str(Path(__file__).resolve().parent.parent / "tests"))
import bench  # noqa: E402
import __graft_entry__ as g  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "bunny"
m = g.load_package()
cfg = bench.CONFIGS[name]
spp = int(sys.argv[2]) if len(sys.argv) > 2 and int(sys.argv[2]) > 0 else cfg["spp"]
mode = m.MODE_B if len(sys.argv) > 3 and sys.argv[3] == "B" else m.MODE_A
opts = set(sys.argv[4:])
flags = (m.RENDER_SMEM_TOP if "smemtop" in opts else 0) | (m.RENDER_POOL_PATHS if "pool" in opts else 0) | (m.RENDER_WAVEFRONT if "wavefront" in opts else 0)
hs = m.HostScene.build(*cfg["scene"], device_tree=name == "stress")
ds = m.DeviceScene(hs)
cam = m.camera_init(bench.camera_for(m, cfg, None, spp))
if "batch" in opts:
    import helpers
    rays = helpers.camera_rays(m, cam, 2_000_000, 3)
    for rep in range(2):
        hits = ds.intersect(rays)
    print(name, "intersect_batch", len(rays), "rays, hit fraction", float((hits["prim_id"] >= 0).mean()), flush=True)
elif "adaptive" in opts:
    for rep in range(2):
        s, q, c, st = ds.render_adaptive(cam, max_spp=spp, max_depth=cfg["depth"], mode=m.MODE_B)
    print(name, "adaptive", {k: st[k] for k in ("samples", "rays", "launches", "ms_total")}, "mean count", float(c.mean()), flush=True)
else:
    for rep in range(2):
        img, st = ds.render(cam, spp=spp, max_depth=cfg["depth"], mode=mode, flags=flags)
        print(name, "rep", rep, {k: st[k] for k in ("samples", "rays", "iterations", "launches", "ms_total")},
              "Msamples/s %.1f" % (st["samples"] / st["ms_total"] / 1e3), flush=True)
