"""Oracle parity AT FULL SIZE (task 1): the BASELINE.json configurations compared with the unmodified
reference compiled into oracle/_ref, not only checked through properties.

  * C5 stress (4 001 620 triangles + 1 000 000 spheres): >= 1 M primary + >= 1 M bounce rays against the reference's
    geom::Bvh::Hit (src/geom/bvh.h:71-119) on BOTH device trees (host SAH tree compiled to pairs, and the linear BVH
    built on the GPU) - ids bit-exact, t within 1e-5 relative (and bit-equal); a band of the 3840x2160 image against
    Camera::GetPixel (src/scene/camera.h:148-174).
  * C1-C4: a >= 64-row band of the full-size image at the configuration's spp, rendered by the reference's pixel loop
    (ora_render_a(y0, y1)), against the same rows of the device image: per pixel-channel 3-sigma test and RMSE bound.

The 3-sigma bar.  For Gaussian pixel means 99.73 % of the channels lie within 3 sigma; the per-pixel sample distributions of
a path tracer are not Gaussian (a pixel that sees a light through two bounces is a rare-event sum) and sigma itself is
estimated from the reference's samples, so the reference does not reach 99.7 % against ITSELF: two reference renders with
different seeds agree on 99.58-99.80 % of the channels of the golden images (tests/golden/image_*.npz: mean vs mean_256).
The bar is therefore the reference's own fraction on the same statistic minus 0.3 points, measured in the same test, and the
fractions are written to gpurun_out/parity_fractions.json (DESIGN.md section 5 lists them).  The only absolute term is
1e-4 * max(ref, 1): float32 accumulation of identical samples (sky pixels have zero variance in the reference's doubles)."""
import json
import os
from pathlib import Path

import numpy as np
import pytest

import bench
import helpers

pytestmark = pytest.mark.gpu

ROOT = Path(__file__).resolve().parents[1]
FLOOR = 1e-4


def _record(key, value):
    out = ROOT / "gpurun_out"
    try:
        out.mkdir(exist_ok=True)
        f = out / "parity_fractions.json"
        d = json.loads(f.read_text()) if f.exists() else {}
        d[key] = value
        f.write_text(json.dumps(d, indent=1, sort_keys=True))
    except OSError:
        pass


def _within(mu, ref, var, n_a, n_b):
    sigma = np.sqrt(var / n_a + var / n_b)
    return np.abs(mu - ref) <= 3 * sigma + FLOOR * np.maximum(ref, 1.0)


def band_parity(rt, name, ds, ora, cam_cfg, cam, spp, depth, y0, y1, gpu_img):
    """gpu_img: rgb_sum of the full image at `spp`.  Two reference renders of rows [y0, y1) with different seeds."""
    cores = len(os.sched_getaffinity(0))
    ref1, m2, _, rays1 = ora.render_a(cam_cfg.as_oracle_cfg(), spp, depth, seed=11, threads=cores, y0=y0, y1=y1)
    ref2, _, _, _ = ora.render_a(cam_cfg.as_oracle_cfg(), spp, depth, seed=12, threads=cores, y0=y0, y1=y1, want_m2=False)
    ref1, ref2, var = ref1[y0:y1], ref2[y0:y1], m2[y0:y1] / max(spp - 1, 1)
    mu = gpu_img[y0:y1].astype(np.float64) / spp
    frac_gpu = float(_within(mu, ref1, var, spp, spp).mean())
    frac_ref = float(_within(ref2, ref1, var, spp, spp).mean())
    rmse_gpu = float(np.sqrt(np.mean((mu - ref1) ** 2)))
    rmse_ref = float(np.sqrt(np.mean((ref2 - ref1) ** 2)))
    _record(f"band:{name}", {"rows": [y0, y1], "spp": spp, "frac_gpu_vs_ref": frac_gpu, "frac_ref_vs_ref": frac_ref,
                             "rmse_gpu_vs_ref": rmse_gpu, "rmse_ref_vs_ref": rmse_ref, "mean_gpu": float(mu.mean()), "mean_ref": float(ref1.mean())})
    print(f"[parity] {name} rows {y0}-{y1} at {spp} spp: 3-sigma fraction gpu-vs-ref {frac_gpu:.5f}, ref-vs-ref {frac_ref:.5f}; "
          f"RMSE {rmse_gpu:.5f} vs {rmse_ref:.5f}")
    assert frac_gpu >= min(0.997, frac_ref - 0.003), f"{name}: {frac_gpu:.5f} of the channels within 3 sigma, the reference against itself {frac_ref:.5f}"
    assert rmse_gpu <= 1.5 * rmse_ref + 1e-4, f"{name}: RMSE {rmse_gpu:.5f} vs reference-vs-reference {rmse_ref:.5f}"
    # no bias: the band-wide mean agrees far tighter than any pixel
    assert abs(mu.mean() - ref1.mean()) <= 0.01 * ref1.mean() + 4 * np.sqrt(2 * np.mean(var) / (spp * mu.size))


@pytest.mark.parametrize("name", ["rtiow", "cornell", "bunny", "earth"])
def test_full_size_band_matches_reference_render(rt, name):
    from oracle import checker
    cfg = bench.CONFIGS[name]
    hs = rt.HostScene.build(*cfg["scene"])
    ora = checker.make_oracle(hs, fast=False)
    if ora.kind != "reference":
        pytest.skip("oracle/_ref was not built (no /root/reference at build time)")
    ds = rt.DeviceScene(hs)
    cam_cfg = bench.camera_for(rt, cfg)
    cam = rt.camera_init(cam_cfg)
    img, st = ds.render(cam, spp=cfg["spp"], max_depth=cfg["depth"])
    assert st["samples"] == cam.width * cam.height * cfg["spp"]
    rows = 64
    y0 = int(cam.height * 0.55) - rows // 2  # the band that crosses the objects and their contact shadows
    band_parity(rt, name, ds, ora, cam_cfg, cam, cfg["spp"], cfg["depth"], y0, y0 + rows, img)
    ds.close()


@pytest.mark.parametrize("name", ["rtiow", "cornell", "bunny", "earth"])
def test_hit_volume_per_config_matches_reference_bvh(rt, name):
    """SURVEY 8c's hit KAT volume: >= 1 M primary rays of the configuration's own camera + >= 1 M secondary rays (two bounce
    generations off the reference's own hit points) per config, ids bit-exact and t bit-equal to geom::Bvh::Hit."""
    from oracle import checker
    cfg = bench.CONFIGS[name]
    hs = rt.HostScene.build(*cfg["scene"])
    ora = checker.make_oracle(hs, fast=False)
    if ora.kind != "reference":
        pytest.skip("oracle/_ref was not built (no /root/reference at build time)")
    ds = rt.DeviceScene(hs)
    cam = rt.camera_init(bench.camera_for(rt, cfg))
    primary = helpers.camera_rays(rt, cam, 1_200_000, 61)
    t0, g0 = ora.intersect(primary)
    assert helpers.check_hits(ds.intersect(primary), t0, g0), f"{name}: primary t not bit-equal"
    b1 = helpers.secondary_rays(rt, primary, t0, g0 >= 0, 62)
    t1, g1 = ora.intersect(b1)
    assert helpers.check_hits(ds.intersect(b1), t1, g1), f"{name}: first-bounce t not bit-equal"
    b2 = helpers.secondary_rays(rt, b1, t1, g1 >= 0, 63)
    t2, g2 = ora.intersect(b2)
    assert helpers.check_hits(ds.intersect(b2), t2, g2), f"{name}: second-bounce t not bit-equal"
    n_sec = len(b1) + len(b2)
    if n_sec < 1_000_000:  # open scenes lose rays to the sky: top the secondary set up with a third generation from more primaries
        more = helpers.camera_rays(rt, cam, 2_000_000, 64)
        tm, gm = ora.intersect(more)
        b3 = helpers.secondary_rays(rt, more, tm, gm >= 0, 65)[: 1_050_000 - n_sec]
        t3, g3 = ora.intersect(b3)
        assert helpers.check_hits(ds.intersect(b3), t3, g3), f"{name}: top-up bounce t not bit-equal"
        n_sec += len(b3)
    assert len(primary) >= 1_000_000 and n_sec >= 1_000_000, (len(primary), n_sec)
    _record(f"hits:{name}", {"primary": int(len(primary)), "secondary": int(n_sec), "primary_hit_fraction": float((g0 >= 0).mean()),
                             "ids": "bit-exact", "t": "bit-equal"})
    ds.close()


# ---- C5: the 5 M-primitive stress scene ------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def stress(rt):
    from oracle import checker
    cfg = bench.CONFIGS["stress"]
    hs = rt.HostScene.build(*cfg["scene"])  # host SAH tree: the export stream the oracle is built from
    ora = checker.make_oracle(hs, fast=False)
    if ora.kind != "reference":
        pytest.skip("oracle/_ref was not built (no /root/reference at build time)")
    yield cfg, hs, ora


def test_c5_full_scene_hits_match_reference_bvh(rt, stress):
    cfg, hs, ora = stress
    d = hs.desc
    assert d.n_triangles == 4_001_620 and d.n_spheres == 1_000_000
    cam = rt.camera_init(bench.camera_for(rt, cfg))
    assert (cam.width, cam.height) == (3840, 2160)
    primary = helpers.camera_rays(rt, cam, 1_100_000, 51)
    t_ref, gid_ref = ora.intersect(primary)
    hit = gid_ref >= 0
    assert hit.mean() > 0.3
    bounce = helpers.secondary_rays(rt, primary, t_ref, hit, 52)
    extra = helpers.ray_batch(rt, 1_050_000 - min(len(bounce), 1_000_000), 53, eye=(0.0, 30.0, 0.0), spread=150.0, jitter=60.0)
    bounce = np.concatenate([bounce, extra])  # incoherent rays from all over the field top the batch up to > 1 M
    assert len(primary) >= 1_000_000 and len(bounce) >= 1_000_000
    tb_ref, gb_ref = ora.intersect(bounce)
    trees = {"host SAH tree": rt.DeviceScene(hs), "device LBVH": rt.DeviceScene(rt.HostScene.build(*cfg["scene"], device_tree=True))}
    for label, ds in trees.items():
        h = ds.intersect(primary)
        assert helpers.check_hits(h, t_ref, gid_ref), f"{label}: primary t not bit-equal to the reference"
        h = ds.intersect(bounce)
        assert helpers.check_hits(h, tb_ref, gb_ref), f"{label}: bounce t not bit-equal to the reference"
        ds.close()
    _record("hits:stress", {"primary": int(len(primary)), "bounce": int(len(bounce)), "primary_hit_fraction": float(hit.mean()),
                            "bounce_hit_fraction": float((gb_ref >= 0).mean()), "ids": "bit-exact", "t": "bit-equal", "trees": list(trees)})


def test_c5_band_of_the_4k_image_matches_reference_render(rt, stress):
    cfg, hs, ora = stress
    cam_cfg = bench.camera_for(rt, cfg)
    cam = rt.camera_init(cam_cfg)
    spp = 16  # of 1024: the reference's pixel loop runs at ~1 Msamples/s on this scene
    ds = rt.DeviceScene(rt.HostScene.build(*cfg["scene"], device_tree=True))  # what bench.py --config stress renders with
    y0 = 1200
    img, st = ds.render(cam, spp=spp, max_depth=cfg["depth"])
    band_parity(rt, "stress", ds, ora, cam_cfg, cam, spp, cfg["depth"], y0, y0 + 32, img)
    ds.close()
