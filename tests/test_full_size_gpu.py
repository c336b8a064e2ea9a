"""BASELINE.json configurations C1-C4 at FULL size through the C ABI, checked by size-independent properties:
sample accounting, the sharding invariant (Philox keyed on the global sample index), pipeline and tree independence
of the traced paths, and agreement of the image statistics with the reduced-size golden renders of the reference."""
import numpy as np
import pytest

import bench

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", params=["rtiow", "cornell", "bunny", "earth"])
def full(request, rt):
    cfg = bench.CONFIGS[request.param]
    hs = rt.HostScene.build(*cfg["scene"])
    ds = rt.DeviceScene(hs)
    cam = rt.camera_init(bench.camera_for(rt, cfg))
    img, st = ds.render(cam, spp=cfg["spp"], max_depth=cfg["depth"])
    yield request.param, cfg, ds, cam, img, st
    ds.close()


def test_sample_accounting_and_finite_radiance(rt, full):
    name, cfg, ds, cam, img, st = full
    assert st["samples"] == cam.width * cam.height * cfg["spp"]
    assert st["launches"] == 1, "the default pipeline renders a configuration in one launch"
    assert st["rays"] >= st["samples"] and st["rays"] <= st["samples"] * cfg["depth"]
    assert np.isfinite(img).all() and (img >= 0).all() and img.mean() > 0
    rgb8 = rt.resolve(img, cfg["spp"])
    assert rgb8.shape == img.shape and rgb8.max() <= 255 and rgb8.std() > 1


def test_spp_shards_add_up_to_the_full_render(rt, full):
    """What an N-GPU run does, on one device: rank slices of the global sample indices, summed."""
    name, cfg, ds, cam, img, st = full
    spp = cfg["spp"]
    acc, rays = np.zeros_like(img), 0
    for r in range(4):
        b, e = rt.multigpu.spp_slice(spp, r, 4)
        part, sp = ds.render(cam, spp=e - b, max_depth=cfg["depth"], spp_begin=b)
        acc += part
        rays += sp["rays"]
    assert rays == st["rays"]
    assert np.abs(acc - img).max() <= 3e-5 * img.max()


def test_wavefront_pipeline_and_device_tree_trace_the_same_paths(rt, full):
    name, cfg, ds, cam, img, st = full
    wave, sw = ds.render(cam, spp=cfg["spp"], max_depth=cfg["depth"], flags=rt.RENDER_WAVEFRONT)
    assert sw["rays"] == st["rays"]
    assert np.abs(wave - img).max() <= 3e-5 * img.max()
    scene = cfg["scene"]
    dev = rt.DeviceScene(rt.HostScene.build(*scene, device_tree=True))
    lb, sl = dev.render(cam, spp=cfg["spp"], max_depth=cfg["depth"])
    dev.close()
    # the device-built tree has no reference slot order: an EXACT-t tie between two primitives (a ray into the edge two
    # Cornell walls share: 9 of 694 M rays) is broken by creation order, and that path continues on the other wall
    assert abs(sl["rays"] - st["rays"]) <= 1e-6 * st["rays"]
    differs = np.abs(lb - img).max(-1) > 3e-5 * img.max()
    assert differs.mean() <= 1e-4


def test_rays_per_sample_matches_the_reference_at_reduced_size(rt, golden_dir, full):
    """The golden renders (reference, reduced resolution, different camera crop for some scenes) count closest-hit
    queries per camera sample; the full-size device render must sit in the same regime."""
    name, cfg, ds, cam, img, st = full
    gname = {"earth": "earth_bubbles"}.get(name, name)
    g = np.load(golden_dir / f"image_{gname}.npz")
    rps = st["rays"] / st["samples"]
    assert 0.6 * float(g["rays_per_sample"]) <= rps <= 1.6 * float(g["rays_per_sample"])
