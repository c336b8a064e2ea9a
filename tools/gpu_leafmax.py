import sys, os, subprocess
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
if len(sys.argv) > 1 and sys.argv[1] == "--one":
    sys.path.insert(0, str(ROOT))
    import numpy as np, bench, __graft_entry__ as g
    m = g.load_package()
    for name in sys.argv[2:]:
        cfg = bench.CONFIGS[name]
        hs = m.HostScene.build(*cfg["scene"]); ds = m.DeviceScene(hs); cam = m.camera_init(bench.camera_for(m, cfg))
        out = []
        for label, flags in (("path", 0), ("wavefront", m.RENDER_WAVEFRONT)):
            ts = []
            for rep in range(4):
                img, st = ds.render(cam, spp=cfg["spp"], max_depth=cfg["depth"], flags=flags)
                ts.append(st["ms_total"])
            out.append("%s %.2f ms" % (label, float(np.median(ts[1:]))))
        hs_c = m.HostScene.build(*cfg["scene"], counters=True); ds_c = m.DeviceScene(hs_c)
        _, sc = ds_c.render(cam, spp=2, max_depth=cfg["depth"])
        print("LEAF_MAX=%s COST=%s %-8s %s | per ray: nodes %.2f sph %.2f tri %.2f rect %.2f | checksum %.1f" % (
            os.environ.get("RTB_LEAF_MAX"), os.environ.get("RTB_PRIM_COST"), name, "  ".join(out), sc["node_visits"] / sc["rays"], sc["sphere_tests"] / sc["rays"],
            sc["triangle_tests"] / sc["rays"], sc["rect_tests"] / sc["rays"], float(img.sum())), flush=True)
else:
    for lm in ("4:2", "4:4", "4:8", "4:16", "1:8"):
        r = subprocess.run([sys.executable, __file__, "--one"] + sys.argv[1:], env=dict(os.environ, RTB_LEAF_MAX=lm.split(":")[0], RTB_PRIM_COST=lm.split(":")[1]), capture_output=True, text=True)
        print(r.stdout.strip() or ("FAILED " + r.stderr[-400:]), flush=True)
