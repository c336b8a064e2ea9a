"""rtb_render_multi (one host thread, worker thread per device, peer-access reduce) on 1..N devices.
Weak: every device renders the configuration's spp; strong: the configuration's spp x 8 split over the devices."""
import sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import bench, __graft_entry__ as g
m = g.load_package()
lib = m.load()
name = sys.argv[1] if len(sys.argv) > 1 else "bunny"
cfg = bench.CONFIGS[name]
hs = m.HostScene.build(*cfg["scene"], device_tree=name == "stress")
cam = m.camera_init(bench.camera_for(m, cfg))
base_spp = 8 if name == "stress" else cfg["spp"]  # (stress: 8 of 1024 spp per unit, the full config takes 5.6 s per render)
n_dev = lib.rtb_device_count()
base = None
for n in [k for k in (1, 2, 4, 8) if k <= n_dev]:
    for label, spp in (("weak", base_spp * n), ("strong", base_spp * 8)):
        wall, dev, red = [], [], []
        for rep in range(5):
            t0 = time.perf_counter()
            img, st = m.render_multi(hs, cam, spp=spp, max_depth=cfg["depth"], n_devices=n)
            wall.append(1e3 * (time.perf_counter() - t0))
            dev.append(st["ms_total"]); red.append(st["ms_shade"])
        w, d = float(np.median(wall[1:])), float(np.median(dev[1:]))
        ms = st["samples"] / 1e3
        print(f"{name} {label:6s} N={n}: {spp:5d} spp  device {d:8.2f} ms (reduce {np.median(red[1:]):.3f})  wall {w:8.2f} ms   {ms / d:9.1f} Msamples/s device, {ms / w:9.1f} end to end", flush=True)
