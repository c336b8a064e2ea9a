"""Generates the golden fixtures from the UNMODIFIED reference (oracle/_ref, built from
/root/reference by oracle/Makefile).  Run here, in the build container:

    python tests/golden/make_golden.py

Outputs (committed):  hits_*.npz  (rays, reference t, reference primitive id)
                      bvh_*.npz   (reference Bvh::nodes()/prim_indices(), or their sha256 for big trees)
                      image_*.npz (reference Path-A mean + per-pixel sample variance at high spp)
                      camera_rays.npz (reference Camera::GetRay samples)
The scenes are rebuilt by the product's host API from fixed seeds; each file stores a digest of
the exported primitive stream so a test can tell "scene changed" from "result changed".
"""
import hashlib
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))
import __graft_entry__ as g  # noqa: E402
import helpers  # noqa: E402
from oracle import ref  # noqa: E402

rt = g.load_package()
OUT = Path(__file__).resolve().parent

SCENES = {
    "cornell": lambda: rt.HostScene.build("cornell"),
    "rtiow": lambda: rt.HostScene.build("rtiow", (11, 1)),
    "bunny": lambda: rt.HostScene.build("bunny", (50,)),
    "earth_bubbles": lambda: rt.HostScene.build("earth_bubbles"),
    "mixed": lambda: rt.HostScene.from_records(*helpers.mixed_scene()),
}
CAMERAS = {  # low-resolution versions of the BASELINE configs' presets
    "cornell": dict(aspect_ratio=1.5, image_width=72, vfov=40.0, lookfrom=(5, 5, -3.6), lookat=(5, 5, 5), focus_dist=9.0, max_depth=20),
    "rtiow": dict(aspect_ratio=1.777, image_width=96, vfov=45.0, lookfrom=(0, 2, 5), lookat=(0, 0, 0), defocus_angle=0.05, focus_dist=5.0, max_depth=10),
    "bunny": dict(image_width=96, vfov=100.0, lookfrom=(2, 2, 5), lookat=(0, 0, 0), max_depth=20),
    "earth_bubbles": dict(image_width=96, vfov=30.0, lookfrom=(0, 1, 3), lookat=(0, 0, 0), defocus_angle=0.1, focus_dist=3.0, max_depth=15),
    "mixed": dict(image_width=80, vfov=60.0, lookfrom=(1, 2.5, 7), lookat=(0, 1, 0), max_depth=12),
}
EYES = {"cornell": ((5, 5, -3.6), 1.0, 4.0), "rtiow": ((0, 2, 8), 3.0, 2.0), "bunny": ((0, 2, 8), 3.0, 2.0),
        "earth_bubbles": ((0, 1, 4), 1.5, 2.0), "mixed": ((0, 2, 6), 2.0, 2.0)}
IMAGE_SPP = {"cornell": 2048, "rtiow": 1024, "bunny": 1024, "earth_bubbles": 1024, "mixed": 1024}


def digest(prims):
    return hashlib.sha256(np.ascontiguousarray(prims).tobytes()).hexdigest()


def main():
    only = set(sys.argv[1:])
    for name, make in SCENES.items():
        if only and name not in only:
            continue
        hs = make()
        prims, mats, texs, img = hs.export()
        rs = ref.RefScene(prims, mats, texs, img)
        eye, spread, jitter = EYES[name]
        cam = rt.camera_init(rt.camera_config(**CAMERAS[name]))
        prim = helpers.camera_rays(rt, cam, 3000, 11)
        rnd = helpers.ray_batch(rt, 3000, 12, eye=eye, spread=spread, jitter=jitter)
        t0, g0 = rs.intersect(prim)
        sec = helpers.secondary_rays(rt, prim, t0, g0 >= 0, 13)[:3000]
        rays = np.concatenate([prim, rnd, sec])
        t, gid = rs.intersect(rays)
        t_w, hit_w, _ = rs.intersect_world(rays)
        assert np.array_equal(t, t_w) and np.array_equal(gid >= 0, hit_w), "id-recovering loop != Bvh::Hit"
        np.savez_compressed(OUT / f"hits_{name}.npz", rays=rays.view(np.float32).reshape(-1, 8), t=t, gid=gid, scene=digest(prims))
        print(name, "hits", len(rays), "hit frac", (gid >= 0).mean())
        # BVH structure
        trees = {}
        for k in range(rs.bvh_count()):
            nodes, idx = rs.bvh_dump(k)
            if nodes.shape[0] <= 2000:
                trees[f"nodes{k}"], trees[f"idx{k}"] = nodes, idx
            else:
                trees[f"sha_nodes{k}"] = hashlib.sha256(nodes.tobytes()).hexdigest()
                trees[f"sha_idx{k}"] = hashlib.sha256(idx.tobytes()).hexdigest()
                trees[f"shape{k}"] = np.array(nodes.shape)
        np.savez_compressed(OUT / f"bvh_{name}.npz", count=rs.bvh_count(), scene=digest(prims), **trees)
        # Path-A image: mean and per-pixel sample variance
        cfg = rt.camera_config(**CAMERAS[name])
        spp = IMAGE_SPP[name]
        mean, m2, secs, nrays = rs.render_a(cfg.as_oracle_cfg(), spp, cfg.max_depth, seed=5)
        var = m2 / (spp - 1)
        # a second, independent reference render at the GPU test's spp: the A-vs-A noise floor
        mean_b, _, _, _ = rs.render_a(cfg.as_oracle_cfg(), 256, cfg.max_depth, seed=6, want_m2=False)
        np.savez_compressed(OUT / f"image_{name}.npz", mean=mean.astype(np.float32), var=var.astype(np.float32), spp=spp,
                            mean_256=mean_b.astype(np.float32), rays_per_sample=nrays / (mean.shape[0] * mean.shape[1] * spp),
                            cfg=cfg.as_oracle_cfg(), scene=digest(prims))
        print(name, "image", mean.shape, "spp", spp, "%.1fs" % secs, "rays/sample %.2f" % (nrays / (mean.shape[0] * mean.shape[1] * spp)))
    if not only or "camera" in only:
        out = {}
        for name, kw in CAMERAS.items():
            cfg = rt.camera_config(**kw).as_oracle_cfg()
            rng = np.random.default_rng(3)
            w = int(cfg[1]); h = max(1, int(w / cfg[0]))
            ij = np.stack([rng.integers(0, w, 500), rng.integers(0, h, 500)], axis=1).astype(np.int32)
            out[f"cfg_{name}"], out[f"ij_{name}"], out[f"rays_{name}"] = cfg, ij, ref.camera_rays(cfg, ij, seed=9)
        np.savez_compressed(OUT / "camera_rays.npz", **out)
        print("camera rays done")


if __name__ == "__main__":
    main()
