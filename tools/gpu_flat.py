"""The flat loop for scenes of a handful of primitives (trace_loop Io::kFlat, RTB_TUNE=3; 11 = at 3 CTAs x 80 registers) against the tree
(RTB_TUNE=0): render time, ray counts, image difference, and the closest hits of a ray batch through both compared bit for bit.
  python tools/gpu_flat.py [--configs cornell,earth] [--reps 5]"""
import json, os, sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))


def main():
    args = sys.argv[1:]
    names, reps = ["cornell", "earth"], 5
    if "--configs" in args:
        i = args.index("--configs"); names = args[i + 1].split(","); del args[i:i + 2]
    if "--reps" in args:
        i = args.index("--reps"); reps = int(args[i + 1]); del args[i:i + 2]
    import numpy as np
    import bench
    import helpers
    import __graft_entry__ as g
    m = g.load_package()
    for name in names:
        cfg = bench.CONFIGS[name]
        hs = m.HostScene.build(*cfg["scene"])
        ds = m.DeviceScene(hs)
        cam = m.camera_init(bench.camera_for(m, cfg))
        rays = helpers.camera_rays(m, cam, 400_000, 7)
        os.environ["RTB_TREE_BATCH"] = "1"
        h2 = ds.intersect(rays)
        bounce = helpers.secondary_rays(m, rays, h2["t"], h2["prim_id"] >= 0, 8)
        b2 = ds.intersect(bounce)
        os.environ.pop("RTB_TREE_BATCH", None)
        os.environ["RTB_FLAT_BATCH"] = "1"
        h4, b4 = ds.intersect(rays), ds.intersect(bounce)
        os.environ.pop("RTB_FLAT_BATCH", None)
        same = all(np.array_equal(a[k], b[k]) for a, b in ((h2, h4), (b2, b4)) for k in ("t", "prim_id"))
        d = hs.desc
        print(json.dumps({"config": name, "hits_identical": bool(same), "rays": int(len(rays) + len(bounce)),
                          "prims": int(d.n_spheres + d.n_triangles + d.n_rects)}), flush=True)
        ref_img = None
        for spp in (cfg["spp"], 1):
            for label, tune in (("tree", "0"), ("flat 4x64", "3"), ("flat 3x80", "11")):
                os.environ["RTB_TUNE"] = tune
                ts = []
                for _ in range(reps + 1):
                    img, st = ds.render(cam, spp=spp, max_depth=cfg["depth"])
                    ts.append(st["ms_total"])
                line = {"config": name, "loop": label, "spp": spp, "ms": round(float(np.median(ts[1:])), 3), "min_ms": round(min(ts), 3), "rays": int(st["rays"])}
                if spp == cfg["spp"]:
                    if ref_img is None:
                        ref_img = img.astype(np.float64)
                    else:
                        line["image_rel_diff"] = float(np.abs(img - ref_img).sum() / ref_img.sum())
                print(json.dumps(line), flush=True)
        os.environ.pop("RTB_TUNE", None)
        ds.close()


if __name__ == "__main__":
    main()
