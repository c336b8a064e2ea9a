"""extend launch time vs rays per launch: serial renders at several pool sizes -> fixed cost (tail) and per-ray cost."""
import sys, os
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import bench, __graft_entry__ as g
m = g.load_package()
name = sys.argv[1] if len(sys.argv) > 1 else "bunny"
cfg = bench.CONFIGS[name]
hs = m.HostScene.build(*cfg["scene"]); ds = m.DeviceScene(hs)
cam = m.camera_init(bench.camera_for(m, cfg))
rows = []
for pool in [1 << 18, 1 << 19, 1 << 20, 1 << 21, 1 << 22, 1 << 23]:
    for rep in range(3):
        img, st = ds.render(cam, spp=cfg["spp"], max_depth=cfg["depth"], pool_size=pool, flags=m.RENDER_SERIAL)
    rows.append((pool, st["iterations"], st["ms_extend"], st["ms_shade"], st["ms_generate"], st["ms_total"], st["rays"]))
    print("pool %8d iters %4d extend %.2f ms (%.1f us/launch, %.3f ns/ray) shade %.2f gen %.2f total %.2f" % (
        pool, st["iterations"], st["ms_extend"], 1e3 * st["ms_extend"] / st["iterations"], 1e6 * st["ms_extend"] / st["rays"],
        st["ms_shade"], st["ms_generate"], st["ms_total"]), flush=True)
A = np.array([[r[1], r[6]] for r in rows], float); y = np.array([r[2] for r in rows])
sol = np.linalg.lstsq(A, y, rcond=None)[0]
print("fit: extend_ms = %.1f us x launches + %.4f ns x rays" % (1e3 * sol[0], 1e6 * sol[1]))
for waves in (1, 2, 4, 6, 8):
    os.environ["RTB_WAVES"] = str(waves)
    for pool in (1 << 20, 1 << 21):
        t = []
        for rep in range(4):
            img, st = ds.render(cam, spp=cfg["spp"], max_depth=cfg["depth"], pool_size=pool)
            t.append(st["ms_total"])
        print("waves %d pool %8d total %.2f ms" % (waves, pool, min(t[1:])), flush=True)
