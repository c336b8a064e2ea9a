import json, os, subprocess, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
if len(sys.argv) > 1 and sys.argv[1] == "--one":
    sys.path.insert(0, str(ROOT))
    import numpy as np, bench, __graft_entry__ as g
    m = g.load_package(); name = sys.argv[2]; cfg = bench.CONFIGS[name]
    hs = m.HostScene.build(*cfg["scene"]); ds = m.DeviceScene(hs); cam = m.camera_init(bench.camera_for(m, cfg))
    out = {}
    for label, flags in (("wavefront", m.RENDER_WAVEFRONT), ("fused", 0)):
        ts = []
        for rep in range(5):
            img, st = ds.render(cam, spp=cfg["spp"], max_depth=cfg["depth"], flags=flags)
            ts.append(st["ms_total"])
        out[label] = round(float(np.median(ts[1:])), 2)
    print(os.environ.get("RTB_LIB"), out, flush=True)
else:
    for lib in sys.argv[2:]:
        r = subprocess.run([sys.executable, __file__, "--one", sys.argv[1]], env=dict(os.environ, RTB_LIB=lib), capture_output=True, text=True)
        print(r.stdout.strip() or ("FAILED " + lib + r.stderr[-300:]), flush=True)
