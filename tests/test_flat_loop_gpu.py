"""The flat loop (trace_loop Io::kFlat): scenes whose reference tree is ONE leaf - geom::Bvh::Hit then tests every primitive of the
scene (src/geom/bvh.h:84-116; Cornell: 9 primitives in a leaf the SAH refused to split, earth: 4) - are traversed without a tree,
every lane testing every primitive in lock step.  Same closest hits as the pair tree and as the reference, bit for bit."""
import numpy as np
import pytest

import bench
import helpers

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", ["cornell", "earth"])
def test_flat_loop_hits_equal_tree_and_reference(rt, monkeypatch, name):
    from oracle import checker
    cfg = bench.CONFIGS[name]
    hs = rt.HostScene.build(*cfg["scene"])
    d = hs.desc
    assert d.n_spheres + d.n_triangles + d.n_rects <= 16
    ds = rt.DeviceScene(hs)
    cam = rt.camera_init(bench.camera_for(rt, cfg))
    primary = helpers.camera_rays(rt, cam, 500_000, 31)
    monkeypatch.setenv("RTB_TREE_BATCH", "1")
    tree = ds.intersect(primary)
    bounce = helpers.secondary_rays(rt, primary, tree["t"], tree["prim_id"] >= 0, 32)
    tree_b = ds.intersect(bounce)
    monkeypatch.delenv("RTB_TREE_BATCH")
    flat, flat_b = ds.intersect(primary), ds.intersect(bounce)  # the default for these scenes
    for a, b in ((tree, flat), (tree_b, flat_b)):
        assert np.array_equal(a["prim_id"], b["prim_id"]) and np.array_equal(a["t"], b["t"]) and np.array_equal(a["ref"], b["ref"])
    ora = checker.make_oracle(hs)
    t_ref, g_ref = ora.intersect(bounce[:200_000])
    assert helpers.check_hits(flat_b[:200_000], t_ref, g_ref), "t not bit-equal to the oracle's"
    ds.close()


@pytest.mark.parametrize("name", ["cornell", "earth"])
def test_flat_loop_renders_what_the_tree_renders(rt, monkeypatch, name):
    """Same primitives hit; the float t handed to shading is the filter's value unless a contest forced the exact replay, and the
    two loops meet different contests, so a path in ~1e8 may continue differently: ray counts agree to 1e-6, images to 1e-4."""
    cfg = bench.CONFIGS[name]
    ds = rt.DeviceScene(rt.HostScene.build(*cfg["scene"]))
    cam = rt.camera_init(bench.camera_for(rt, cfg, 240))
    flat, sf = ds.render(cam, spp=32, max_depth=cfg["depth"])
    wave, sw = ds.render(cam, spp=32, max_depth=cfg["depth"], flags=rt.RENDER_WAVEFRONT)  # extend + drain use the flat loop too
    monkeypatch.setenv("RTB_TUNE", "0")
    tree, st = ds.render(cam, spp=32, max_depth=cfg["depth"])
    assert sf["samples"] == st["samples"] == sw["samples"]
    assert abs(sf["rays"] - st["rays"]) <= 1e-6 * st["rays"] + 2 and abs(sw["rays"] - sf["rays"]) <= 1e-6 * st["rays"] + 2
    scale = max(1.0, float(tree.max()))
    assert np.abs(flat - tree).sum() <= 1e-4 * np.abs(tree).sum() and np.abs(wave - flat).sum() <= 1e-4 * np.abs(tree).sum()
    assert np.abs(flat - tree).max() <= 0.05 * scale * 32  # a diverged path changes a pixel by at most a few samples' radiance
    ds.close()


def test_large_scenes_keep_the_tree(rt, monkeypatch):
    """RTB_TUNE=3 on a scene of 69 452 triangles is ignored (the flat loop is refused beyond 4096 primitives)."""
    ds = rt.DeviceScene(rt.HostScene.build("bunny", (50,)))
    cam = rt.camera_init(rt.camera_config(aspect_ratio=1.5, image_width=96, samples_per_pixel=4, max_depth=8, vfov=30.0, lookfrom=(0, 3, 8),
                                          lookat=(0, 0.8, 0), focus_dist=8.0))
    a, sa = ds.render(cam, spp=4, max_depth=8)
    monkeypatch.setenv("RTB_TUNE", "3")
    b, sb = ds.render(cam, spp=4, max_depth=8)
    assert sa["rays"] == sb["rays"] and sb["ms_total"] < 1000.0
    ds.close()
