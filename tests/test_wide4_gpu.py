"""RTB_RENDER_WIDE4 (the 4-wide traversal loop, trace_loop node_step4; an opt-in experiment): the closest hits are
those of the pair loop and of the reference's geom::Bvh::Hit (src/geom/bvh.h:71-119) bit for bit, the render traces the same paths."""
import numpy as np
import pytest

import helpers

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name,params", [("bunny", (50,)), ("rtiow", (11, 1)), ("cornell", ()), ("earth_bubbles", ())])
def test_quad_loop_hits_equal_pair_loop_and_reference(rt, monkeypatch, name, params):
    from oracle import checker
    monkeypatch.setenv("RTB_WIDE4", "1")
    hs = rt.HostScene.build(name, params)
    assert hs.tree_info()["quads"] > 0
    ds = rt.DeviceScene(hs)
    rays = helpers.ray_batch(rt, 300_000, 91, eye=(0.0, 2.0, 6.0), spread=4.0, jitter=3.0)
    monkeypatch.delenv("RTB_WIDE4_BATCH", raising=False)
    pairs = ds.intersect(rays)
    monkeypatch.setenv("RTB_WIDE4_BATCH", "1")
    quads = ds.intersect(rays)
    assert np.array_equal(pairs["prim_id"], quads["prim_id"]) and np.array_equal(pairs["t"], quads["t"])
    assert 0.02 < (quads["prim_id"] >= 0).mean()
    ora = checker.make_oracle(hs)
    t_ref, gid_ref = ora.intersect(rays[:100_000])
    assert helpers.check_hits(quads[:100_000], t_ref, gid_ref), "t not bit-equal to the oracle's"
    ds.close()


def test_quad_loop_renders_the_same_paths(rt, monkeypatch):
    monkeypatch.setenv("RTB_WIDE4", "1")
    ds = rt.DeviceScene(rt.HostScene.build("bunny", (50,)))
    cfg = rt.camera_config(aspect_ratio=1.5, image_width=240, samples_per_pixel=32, max_depth=20, vfov=30.0, lookfrom=(0, 3, 8),
                           lookat=(0, 0.8, 0), focus_dist=8.0)
    cam = rt.camera_init(cfg)
    a, sa = ds.render(cam, spp=32, max_depth=20)
    b, sb = ds.render(cam, spp=32, max_depth=20, flags=rt.RENDER_WIDE4)
    assert sa["rays"] == sb["rays"] and sa["samples"] == sb["samples"]  # same paths: same closest hits, same Philox streams
    assert np.abs(a - b).max() <= 1e-3 * max(1.0, float(a.max()))       # float atomics in another order
    ds.close()


def test_wide4_is_refused_without_quads(rt, monkeypatch):
    monkeypatch.delenv("RTB_WIDE4", raising=False)
    ds = rt.DeviceScene(rt.HostScene.build("cornell"))
    cam = rt.camera_init(rt.camera_config(aspect_ratio=1.0, image_width=32, samples_per_pixel=1, max_depth=4, vfov=40.0, lookfrom=(5, 5, -3.6),
                                          lookat=(5, 5, 5), focus_dist=9.0))
    with pytest.raises(rt.RtbError) as e:
        ds.render(cam, spp=1, max_depth=4, flags=rt.RENDER_WIDE4)
    assert e.value.code == rt.RTB_EINVAL and "RTB_WIDE4" in str(e.value)
    ds.close()
