"""The 4-wide tree experiment (rtb_compile.h LayoutWide, trace_loop node_step4): C1-C4 rendered with the pair loop and
with the quad loop (4 CTAs x 64 registers, and 3 x 80), at the configuration's spp and at 1 spp (the launch floor), the images and ray
counts compared, and the closest hits of a ray batch through both trees compared bit for bit.
  python tools/gpu_wide4.py [--configs bunny,rtiow,cornell,earth] [--reps 5] [--spp 12,13: these sample counts instead (what one of 8 GPUs
  renders in the spp split of section 6), no hit comparison]
With RTB_LIB=librtb200_counters.so the per-ray step counts are printed as well (timings are then meaningless)."""
import json, os, sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))


def main():
    args = sys.argv[1:]
    names, reps = ["bunny", "rtiow", "cornell", "earth"], 5
    if "--configs" in args:
        i = args.index("--configs"); names = args[i + 1].split(","); del args[i:i + 2]
    spp_list = None
    if "--spp" in args:
        i = args.index("--spp"); spp_list = [int(v) for v in args[i + 1].split(",")]; del args[i:i + 2]
    if "--reps" in args:
        i = args.index("--reps"); reps = int(args[i + 1]); del args[i:i + 2]
    os.environ["RTB_WIDE4"] = "1"  # read when the tree is compiled
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
        # closest hits: the pair loop and the quad loop must agree bit for bit (same exact tests, same tie rule)
        rays = helpers.camera_rays(m, cam, 400_000, 7)
        os.environ.pop("RTB_WIDE4_BATCH", None)
        h2 = ds.intersect(rays)
        bounce = helpers.secondary_rays(m, rays, h2["t"], h2["prim_id"] >= 0, 8)
        b2 = ds.intersect(bounce)
        os.environ["RTB_WIDE4_BATCH"] = "1"
        h4, b4 = ds.intersect(rays), ds.intersect(bounce)
        os.environ.pop("RTB_WIDE4_BATCH", None)
        same = all(np.array_equal(a[k], b[k]) for a, b in ((h2, h4), (b2, b4)) for k in ("t", "prim_id"))
        print(json.dumps({"config": name, "hits_identical": bool(same), "rays": int(len(rays) + len(bounce)),
                          "hit_fraction": float((h2["prim_id"] >= 0).mean()), "tree": hs.tree_info()}), flush=True)
        ref_img = None
        for spp in spp_list or (cfg["spp"], 1):
            for label, tune in (("pairs", None), ("quads 4x64", "4"), ("quads 3x80", "12")):
                os.environ.pop("RTB_TUNE", None)
                if tune:
                    os.environ["RTB_TUNE"] = tune
                ts = []
                for _ in range(reps + 1):
                    img, st = ds.render(cam, spp=spp, max_depth=cfg["depth"])
                    ts.append(st["ms_total"])
                line = {"config": name, "tree": label, "spp": spp, "ms": round(float(np.median(ts[1:])), 3), "min_ms": round(min(ts), 3), "rays": int(st["rays"])}
                if st.get("node_visits"):
                    line["box_tests_per_ray"] = round(st["node_visits"] / st["rays"], 2)
                if spp == cfg["spp"]:
                    if ref_img is None:
                        ref_img = img.astype(np.float64)
                    else:  # same paths: only the order of the float atomics differs
                        line["image_rel_diff"] = float(np.abs(img - ref_img).sum() / ref_img.sum())
                print(json.dumps(line), flush=True)
        os.environ.pop("RTB_TUNE", None)
        ds.close()


if __name__ == "__main__":
    main()
