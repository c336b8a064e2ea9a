"""Pins oracle/port.c (the plain-C restatement of Aabb/Sphere/Triangle/rect::Hit and the
id-carrying Bvh::Hit loop) against outputs of the reference itself: tests/golden/hits_*.npz were
produced by the unmodified reference compiled from /root/reference (tests/golden/make_golden.py)."""
import numpy as np
import pytest

from test_host_scene import SCENES


@pytest.mark.parametrize("name", list(SCENES))
def test_port_matches_reference_golden_hits(rt, golden_dir, name):
    from oracle import port
    hs = SCENES[name](rt)
    ora = port.PortOracle(hs)
    g = np.load(golden_dir / f"hits_{name}.npz")
    rays = np.ascontiguousarray(g["rays"]).view(rt.RAY_DTYPE).reshape(-1)
    t, gid = ora.intersect(rays)
    assert np.array_equal(gid, g["gid"]), "primitive ids differ from the reference"
    hit = g["gid"] >= 0
    assert np.array_equal(t[hit], g["t"][hit]), "t not bit-equal to the reference"
    assert hit.any() and (~hit).any()


def test_port_matches_compiled_reference_on_fresh_rays(rt):
    """Where oracle/_ref exists (this container, and it travels to the GPU box): a fresh seeded batch."""
    from oracle import port, ref
    if not ref.available():
        pytest.skip("oracle/_ref not built")
    from oracle import checker
    import helpers
    hs = SCENES["mixed"](rt)
    rays = helpers.ray_batch(rt, 20000, 5, eye=(0, 1, 4), spread=3.0, jitter=2.0)
    t_ref, g_ref = checker.make_oracle(hs).intersect(rays)
    t, g = port.PortOracle(hs).intersect(rays)
    assert np.array_equal(g, g_ref) and np.array_equal(t[g_ref >= 0], t_ref[g_ref >= 0])
