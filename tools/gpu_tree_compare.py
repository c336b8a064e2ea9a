"""Device trees compared on the BASELINE configs: reference topology + leaf split (default), full SAH rebuild
(RTB_TREE_REBUILD=1), device-built LBVH.  Render time from the product build, visits per ray from the counters build."""
import os, subprocess, sys, json
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
if len(sys.argv) > 1 and sys.argv[1] == "--one":
    sys.path.insert(0, str(ROOT))
    import numpy as np
    import bench, __graft_entry__ as g
    m = g.load_package()
    name, dev = sys.argv[2], sys.argv[3] == "1"
    cfg = bench.CONFIGS[name]
    spp = 4 if name == "stress" else cfg["spp"]
    cam = m.camera_init(bench.camera_for(m, cfg))
    ds = m.DeviceScene(m.HostScene.build(*cfg["scene"], device_tree=dev))
    ts = []
    for rep in range(5):
        img, st = ds.render(cam, spp=spp, max_depth=cfg["depth"])
        ts.append(st["ms_total"])
    dc = m.DeviceScene(m.HostScene.build(*cfg["scene"], device_tree=dev, counters=True))
    _, sc = dc.render(cam, spp=min(spp, 4), max_depth=cfg["depth"])
    print(json.dumps({"ms": float(np.median(ts[1:])), "nodes": sc["node_visits"] / sc["rays"], "tri": sc["triangle_tests"] / sc["rays"],
                      "sph": sc["sphere_tests"] / sc["rays"], "rect": sc["rect_tests"] / sc["rays"], "chk": float(img.sum())}))
else:
    for name in sys.argv[1:] or ["bunny", "rtiow", "cornell", "earth"]:
        for label, env, dev in (("reference topology + leaf split", {"RTB_TREE_REBUILD": "0"}, "0"), ("full SAH rebuild (3 axes)      ", {"RTB_TREE_REBUILD": "1"}, "0"),
                                ("device LBVH                    ", {}, "1")):
            r = subprocess.run([sys.executable, __file__, "--one", name, dev], env=dict(os.environ, **env), capture_output=True, text=True)
            try:
                d = json.loads(r.stdout.strip().splitlines()[-1])
                print(f"{name:8s} {label}: {d['ms']:8.2f} ms   node visits/ray {d['nodes']:6.2f}  tri {d['tri']:.2f} sph {d['sph']:.2f} rect {d['rect']:.2f}  checksum {d['chk']:.1f}", flush=True)
            except Exception:
                print(name, label, "FAILED", r.stderr[-300:], flush=True)
