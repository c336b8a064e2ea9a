"""The 4-wide form of the device tree (rtb_compile.h LayoutWide; an opt-in experiment): host-side structure checks
through rtb_tree_info - no GPU.  The traversal itself is compared with the pair loop and the oracle in tests/test_wide4_gpu.py."""
import pytest

SCENES = [("bunny", (50,)), ("rtiow", (11, 1)), ("cornell", ()), ("earth", ()), ("earth_bubbles", ())]


@pytest.mark.parametrize("name,params", SCENES)
def test_quads_hold_every_primitive_once(rt, monkeypatch, name, params):
    monkeypatch.setenv("RTB_WIDE4", "1")
    hs = rt.HostScene.build(name, params)
    d = hs.desc
    n = d.n_spheres + d.n_triangles + d.n_rects
    info = hs.tree_info()
    # leaves partition the primitives in both forms (a one-primitive scene holds it in both halves of its only record)
    assert info["pair_prims"] == info["quad_prims"] == max(n, 2)
    assert info["pair_box_escapes"] == 0 and info["quad_box_escapes"] == 0
    assert 0 < info["quads"] <= info["pairs"] and 3 * info["quads"] + 1 >= info["pairs"]  # a quad replaces 1-3 pairs
    assert info["quad_stack"] + 2 <= 64  # the 64-entry traversal stack (bvh.h:80) bounds the quad loop too


def test_quads_are_only_built_on_request(rt, monkeypatch):
    monkeypatch.delenv("RTB_WIDE4", raising=False)
    info = rt.HostScene.build("bunny", (50,)).tree_info()
    assert info["quads"] == 0 and info["pairs"] > 60000 and info["pair_prims"] == 69452
