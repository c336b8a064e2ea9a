"""Tuning harness: time one BASELINE config with several builds of the library (make VARIANT=x EXTRA=-D...).
  python tools/gpu_variants.py bunny librtb200.so librtb200_v1.so ...
Each library runs in its own process (RTB_LIB) so that pools and caches do not interact."""
import json, os, subprocess, sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
if len(sys.argv) > 1 and sys.argv[1] == "--one":
    sys.path.insert(0, str(ROOT))
    import numpy as np
    import bench
    import __graft_entry__ as g
    m = g.load_package()
    name = sys.argv[2]
    cfg = bench.CONFIGS[name]
    hs = m.HostScene.build(*cfg["scene"])
    ds = m.DeviceScene(hs)
    cam = m.camera_init(bench.camera_for(m, cfg))
    spp = int(os.environ.get("VAR_SPP", cfg["spp"]))
    tot, ser = [], []
    for rep in range(6):
        img, st = ds.render(cam, spp=spp, max_depth=cfg["depth"])
        if rep >= 2:
            tot.append(st["ms_total"])
    for rep in range(3):
        img, st = ds.render(cam, spp=spp, max_depth=cfg["depth"], flags=m.RENDER_SERIAL)
        ser.append((st["ms_extend"], st["ms_shade"], st["ms_generate"], st["ms_total"]))
    ser = np.mean(ser, axis=0)
    print(json.dumps({"lib": os.environ.get("RTB_LIB", "librtb200.so"), "config": name, "ms_total": float(np.median(tot)),
                      "msamples_s": st["samples"] / float(np.median(tot)) / 1e3, "serial_extend_ms": float(ser[0]),
                      "serial_shade_ms": float(ser[1]), "serial_gen_ms": float(ser[2]), "serial_total_ms": float(ser[3]),
                      "checksum": float(img.sum())}), flush=True)
else:
    name = sys.argv[1]
    for lib in sys.argv[2:]:
        env = dict(os.environ, RTB_LIB=lib)
        r = subprocess.run([sys.executable, __file__, "--one", name], env=env, capture_output=True, text=True)
        print(r.stdout.strip() or ("FAILED " + lib + ": " + r.stderr[-400:]), flush=True)
