"""A/B harness for the render kernels: every BASELINE config (stress at reduced spp) with the default path kernel (AB_POOL=1: and the pool
kernel; AB_MODE_B=1: both estimators), for each library given (make VARIANT=x EXTRA=-D... builds librtb200_x.so).
  python tools/gpu_ab.py [--configs bunny,rtiow,...] [--stress-spp 4] [librtb200.so librtb200_v1.so ...]
Each library runs in its own process (RTB_LIB).  Prints one JSON line per (library, config, kernel)."""
import json, os, subprocess, sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent


def one(names, stress_spp):
    sys.path.insert(0, str(ROOT))
    import numpy as np
    import bench
    import __graft_entry__ as g
    m = g.load_package()
    for name in names:
        cfg = bench.CONFIGS[name]
        if os.environ.get("AB_COST_ORDER"):  # the default kernel with and without the cost-ordered schedule, same library
            ds = m.DeviceScene(m.HostScene.build(*cfg["scene"], device_tree=name == "stress"))
            cam = m.camera_init(bench.camera_for(m, cfg))
            spp = stress_spp if name == "stress" else cfg["spp"]
            for label, env in (("cost_order", "1"), ("plain", None)):
                os.environ.pop("RTB_COST_ORDER", None)
                if env:
                    os.environ["RTB_COST_ORDER"] = env
                for div in (1, 8):  # the configuration's spp, and 1/8 of it (what one of 8 GPUs renders in the spp split)
                    ts = []
                    for rep in range(6):
                        img, st = ds.render(cam, spp=max(1, spp // div), max_depth=cfg["depth"])
                        ts.append(st["ms_total"])
                    print(json.dumps({"config": name, "schedule": label, "spp": max(1, spp // div), "ms": round(float(np.median(ts[1:])), 3), "min_ms": round(min(ts), 3),
                                      "rays": st["rays"], "launches": st["launches"], "checksum": float(img.astype(np.float64).sum())}), flush=True)
            os.environ.pop("RTB_COST_ORDER", None)
            ds.close()
            continue
        dev_tree = name == "stress"
        ds = m.DeviceScene(m.HostScene.build(*cfg["scene"], device_tree=dev_tree))
        cam = m.camera_init(bench.camera_for(m, cfg))
        spp = stress_spp if name == "stress" else cfg["spp"]
        imgs = {}
        for label, flags in (("path", 0),) + ((("pool", m.RENDER_POOL_PATHS),) if os.environ.get("AB_POOL") else ()):
            for mode, mname in ((m.MODE_A, "A"),) + (((m.MODE_B, "B"),) if os.environ.get("AB_MODE_B") else ()):
                ts = []
                for rep in range(5):
                    img, st = ds.render(cam, spp=spp, max_depth=cfg["depth"], flags=flags, mode=mode)
                    ts.append(st["ms_total"])
                t = float(np.median(ts[1:]))
                imgs[(label, mname)] = img
                print(json.dumps({"lib": os.environ.get("RTB_LIB", "librtb200.so"), "config": name, "kernel": label, "mode": mname, "ms": round(t, 3),
                                  "min_ms": round(min(ts), 3), "msamples_s": round(st["samples"] / t / 1e3, 1), "mrays_s": round(st["rays"] / t / 1e3, 1),
                                  "rays": st["rays"]}), flush=True)
        for mname in ("A", "B"):
            if ("path", mname) in imgs:
                print(json.dumps({"lib": os.environ.get("RTB_LIB", "librtb200.so"), "config": name, "mode": mname,
                                  "image_checksum": float(imgs[("path", mname)].astype(np.float64).sum())}), flush=True)
        ds.close()


if __name__ == "__main__":
    args = sys.argv[1:]
    names, stress_spp = ["rtiow", "cornell", "bunny", "earth", "stress"], 4
    if "--configs" in args:
        i = args.index("--configs"); names = args[i + 1].split(","); del args[i:i + 2]
    if "--stress-spp" in args:
        i = args.index("--stress-spp"); stress_spp = int(args[i + 1]); del args[i:i + 2]
    if args and args[0] == "--one":
        one(names, stress_spp)
    else:
        for lib in args or ["librtb200.so"]:
            r = subprocess.run([sys.executable, __file__, "--configs", ",".join(names), "--stress-spp", str(stress_spp), "--one"],
                               env=dict(os.environ, RTB_LIB=lib), capture_output=True, text=True)
            print(r.stdout.strip() or ("FAILED " + lib + ": " + r.stderr[-600:]), flush=True)
            if r.returncode:
                print("FAILED " + lib + ": " + r.stderr[-600:], flush=True)
