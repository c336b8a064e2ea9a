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


def _pixel_state_golden(golden_dir):
    from golden.make_golden_pixel_state import sequences
    g = np.load(golden_dir / "pixel_state.npz")
    x = sequences(int(g["n_seq"]), int(g["len"]), int(g["seed"]))
    assert float(x.astype(np.float64).sum()) == float(g["checksum"]), "sample generator changed: regenerate the fixture"
    return g, x


def test_port_stop_rule_matches_reference_pixel_state(golden_dir):
    """oracle/port.c port_pixel_state_run against the reference's own RecordSample / IsConverged
    (tests/golden/pixel_state.npz): stop counts, PixelState::sum and ::m2 bit-equal."""
    from oracle import port
    g, x = _pixel_state_golden(golden_dir)
    for i, (thr, mn) in enumerate(g["settings"]):
        stop, s, m2 = port.pixel_state_run(x.astype(np.float64), float(thr), int(mn))
        assert np.array_equal(stop, g[f"stop_{i}"])
        assert np.array_equal(s, g[f"sum_{i}"]) and np.array_equal(m2, g[f"m2_{i}"])
        assert (stop < x.shape[1]).any() and (stop == x.shape[1]).any() and stop.min() >= int(mn)
