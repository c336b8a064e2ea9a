"""Where the flat loop and the tree disagree, and who is right.  Renders a configuration through the wavefront pipeline with
RTB_DEBUG_FLAT_VS_TREE=1 (every queue of rays goes through BOTH loops, disagreeing rays are printed by the library), then hands
those rays to the reference (oracle/_ref) and to rtb_intersect_batch with either loop.
  python tools/gpu_flat_mismatch.py [config] [spp]"""
import json, os, subprocess, sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

name = sys.argv[1] if len(sys.argv) > 1 else "cornell"
if len(sys.argv) > 3 and sys.argv[3] == "--render":  # child: the render whose stderr carries the mismatch lines
    import bench
    import __graft_entry__ as g
    m = g.load_package()
    cfg = bench.CONFIGS[name]
    ds = m.DeviceScene(m.HostScene.build(*cfg["scene"]))
    cam = m.camera_init(bench.camera_for(m, cfg))
    img, st = ds.render(cam, spp=int(sys.argv[2]), max_depth=cfg["depth"], flags=m.RENDER_WAVEFRONT)
    print(json.dumps({"rays": int(st["rays"])}))
    sys.exit(0)

import numpy as np
import bench
import __graft_entry__ as g
from oracle import checker

m = g.load_package()
cfg = bench.CONFIGS[name]
spp = int(sys.argv[2]) if len(sys.argv) > 2 else cfg["spp"]
r = subprocess.run([sys.executable, __file__, name, str(spp), "--render"], capture_output=True, text=True, env=dict(os.environ, RTB_DEBUG_FLAT_VS_TREE="1"))
lines = [json.loads(ln.replace(" inf", " Infinity").replace(" -inf", " -Infinity").replace(" nan", " NaN").replace(" -nan", " NaN"))
         for ln in r.stderr.splitlines() if ln.startswith("{")]
summary = [ln for ln in lines if "flat_vs_tree_mismatches" in ln]
rays = [ln for ln in lines if "mismatch_ray" in ln]
print(json.dumps({"config": name, "spp": spp, "render": r.stdout.strip(), "library": summary}), flush=True)
if rays:
    hs = m.HostScene.build(*cfg["scene"])
    ds = m.DeviceScene(hs)
    ora = checker.make_oracle(hs, fast=False)
    batch = np.zeros(len(rays), m.RAY_DTYPE)
    for k, ln in enumerate(rays):
        v = ln["mismatch_ray"]
        batch["o"][k], batch["tmin"][k], batch["d"][k], batch["tmax"][k] = v[0:3], v[3], v[4:7], v[7]
    os.environ["RTB_TREE_BATCH"] = "1"
    tree = ds.intersect(batch)
    os.environ.pop("RTB_TREE_BATCH", None)
    os.environ["RTB_FLAT_BATCH"] = "1"
    flat = ds.intersect(batch)
    t_ref, g_ref = ora.intersect(batch)
    for k in range(len(rays)):
        print(json.dumps({"ray": rays[k]["mismatch_ray"], "tree": [int(tree["prim_id"][k]), float(tree["t"][k])],
                          "flat": [int(flat["prim_id"][k]), float(flat["t"][k])], "reference": [int(g_ref[k]), float(t_ref[k])]}), flush=True)
    print(json.dumps({"reference_agrees_with_flat": int(sum(int(g_ref[k]) == int(flat["prim_id"][k]) and (g_ref[k] < 0 or t_ref[k] == flat["t"][k]) for k in range(len(rays)))),
                      "reference_agrees_with_tree": int(sum(int(g_ref[k]) == int(tree["prim_id"][k]) and (g_ref[k] < 0 or t_ref[k] == tree["t"][k]) for k in range(len(rays)))),
                      "of": len(rays)}))
