"""Estimator B with the reference's per-pixel adaptive stop: rtb_render_adaptive against
integrator::PixelState / IsConverged (src/integrator/pixel_state.h:13-72) as WavefrontRenderer::Render applies
them (src/renderer/wavefront.cc:42-43,62-82,120-205,229-241), and against the golden runs of the UNMODIFIED
reference renderer (tests/golden/pathb_*.npz, made by tests/golden/make_golden_pathb.py)."""
import numpy as np
import pytest

from test_host_scene import SCENES

pytestmark = pytest.mark.gpu

PATHB_CAMERAS = {  # = tests/golden/make_golden.py CAMERAS (the fixtures were rendered with these)
    "cornell": dict(aspect_ratio=1.5, image_width=72, vfov=40.0, lookfrom=(5, 5, -3.6), lookat=(5, 5, 5), focus_dist=9.0, max_depth=20),
    "rtiow": dict(aspect_ratio=1.777, image_width=96, vfov=45.0, lookfrom=(0, 2, 5), lookat=(0, 0, 0), defocus_angle=0.05, focus_dist=5.0, max_depth=10),
    "earth_bubbles": dict(image_width=96, vfov=30.0, lookfrom=(0, 1, 3), lookat=(0, 0, 0), defocus_angle=0.1, focus_dist=3.0, max_depth=15),
    "mixed": dict(image_width=80, vfov=60.0, lookfrom=(1, 2.5, 7), lookat=(0, 1, 0), max_depth=12),
}


@pytest.fixture(scope="module")
def scenes(rt):
    return {name: rt.DeviceScene(make(rt)) for name, make in SCENES.items() if name in PATHB_CAMERAS}


def is_converged(sum_, sumsq, n, thr, min_spp):
    """integrator::IsConverged (pixel_state.h:53-72) in float64 from the sum / sum-of-squares planes."""
    n = n.astype(np.float64)[..., None]
    mean = sum_ / np.maximum(n, 1)
    var = np.maximum(sumsq - sum_ * mean, 0.0) / np.maximum(n - 1, 1)
    err = np.sqrt(var) / np.sqrt(np.maximum(n, 1))
    ok = (err / np.maximum(np.abs(mean), 1e-3) <= thr).all(-1)
    return ok & (n[..., 0] >= min_spp)


def test_never_converging_equals_fixed_spp(rt, scenes):
    """min_samples = max: every pixel takes all its samples, and they are the SAME samples (Philox on the
    global sample index) the fixed-spp render traces."""
    ds = scenes["mixed"]
    cam = rt.camera_init(rt.camera_config(**PATHB_CAMERAS["mixed"]))
    fixed, sf = ds.render(cam, spp=24, max_depth=12, mode=rt.MODE_B)
    s, q, c, st = ds.render_adaptive(cam, max_spp=24, max_depth=12, min_samples=24, rel_threshold=0.0)
    assert (c == 24).all()
    assert st["samples"] == sf["samples"] == 24 * cam.width * cam.height
    assert st["rays"] == sf["rays"]
    assert np.abs(s - fixed).max() <= 2e-5 * fixed.max()
    assert (q >= 0).all() and (q.sum(-1) > 0).any()


def test_stop_rule_replayed_sample_by_sample(rt, scenes):
    """A pixel stops at n exactly when IsConverged first holds at n (n >= kMinSamples): replay the decision from the
    planes of runs capped at every k."""
    ds = scenes["mixed"]
    cam = rt.camera_init(rt.camera_config(**PATHB_CAMERAS["mixed"]))
    thr, mn, top = 0.05, 16, 40
    runs = {k: ds.render_adaptive(cam, max_spp=k, max_depth=12, rel_threshold=thr, min_samples=mn) for k in range(mn, top + 1)}
    final = runs[top][2]
    assert final.min() >= mn and final.max() <= top
    assert (final < top).mean() > 0.1, "scene too noisy for this test: nothing converged"
    borderline = 0
    for k in range(mn, top + 1):
        s, q, c, _ = runs[k]
        assert np.array_equal(c, np.minimum(final, k)), f"cap {k}: counts are not the prefix of the longer run"
        at_k = c == k
        conv = is_converged(s.astype(np.float64), q.astype(np.float64), c, thr, mn)
        # float planes against a float64 replay: allow the decision to differ within 1e-3 of the threshold
        hi = is_converged(s.astype(np.float64), q.astype(np.float64), c, thr * 1.001, mn)
        lo = is_converged(s.astype(np.float64), q.astype(np.float64), c, thr * 0.999, mn)
        stopped_here = at_k & (final == k)
        went_on = at_k & (final > k)
        if k < top:
            assert not (stopped_here & ~hi).any(), f"n={k}: a pixel stopped although the rule does not hold"
        assert not (went_on & lo).any(), f"n={k}: a pixel went on although the rule holds"
        borderline += int((hi != lo).sum())
        del conv
    assert borderline < 0.01 * final.size * (top - mn)


def test_luminance_criterion(rt, scenes):
    """RTB_ADAPT_LUMINANCE: integrator::AdaptiveSampler's test (sampler.h:61-71) on estimator A."""
    ds = scenes["rtiow"]
    cam = rt.camera_init(rt.camera_config(**PATHB_CAMERAS["rtiow"]))
    thr, mn, mx = 0.03, 8, 64
    s, q, c, st = ds.render_adaptive(cam, max_spp=mx, max_depth=10, rel_threshold=thr, min_samples=mn, criterion=rt.ADAPT_LUMINANCE, mode=rt.MODE_A)
    # samples are traced speculatively in chunks of at most 64; what the stop rule did not need is discarded, not recorded
    assert c.min() >= mn and c.max() <= mx and int(c.sum()) <= st["samples"] <= int(c.sum()) + 63 * c.size
    lum = np.array([0.2126, 0.7152, 0.0722])
    n = c.astype(np.float64)[..., None]
    mean = s / n
    var = q / n - mean * mean
    err = np.sqrt(np.maximum((var * lum).sum(-1), 0) / n[..., 0])
    rel = err / ((mean * lum).sum(-1) + 1e-3)
    stopped = c < mx
    assert 0.05 < stopped.mean() < 1.0
    assert (rel[stopped] < thr * 1.001).all()
    # sky pixels are nearly noise-free and stop at the minimum; the glass / metal spheres keep sampling
    assert (c == mn).mean() > 0.05


@pytest.mark.parametrize("name", list(PATHB_CAMERAS))
def test_matches_reference_wavefront_renderer(rt, golden_dir, scenes, name):
    """Against RUNS runs of the reference's own WavefrontRenderer: (i) the number of pixels still sampling in every
    pass (the reference logs it as "queue="), (ii) the 8-bit image."""
    g = np.load(golden_dir / f"pathb_{name}.npz")
    max_samples, depth, runs = int(g["max_samples"]), int(g["max_depth"]), int(g["runs"])
    cam = rt.camera_init(rt.camera_config(**PATHB_CAMERAS[name]))
    n_pix = cam.width * cam.height
    s, q, c, st = scenes[name].render_adaptive(cam, max_spp=max_samples, max_depth=depth, rel_threshold=0.05, min_samples=16, mode=rt.MODE_B)
    assert int(c.sum()) <= st["samples"] <= int(c.sum()) + 63 * c.size
    active = np.array([(c > k).sum() for k in range(max_samples)], dtype=np.float64)
    ref_a, ref_s = g["active_mean"], g["active_std"]
    assert np.all(active[:16] == n_pix) and np.all(ref_a[:16] == n_pix)
    tol = 5 * ref_s * np.sqrt(1 + 1 / runs) + 0.02 * n_pix
    assert np.all(np.abs(active - ref_a) <= tol), f"{name}: active-pixel curve off by {np.abs(active - ref_a).max():.0f} pixels"
    assert abs(c.sum() - ref_a.sum()) <= 0.03 * ref_a.sum(), f"{name}: total samples {c.sum()} vs reference {ref_a.sum():.0f}"
    # image: the device's bytes against the mean of the reference's runs, in units of the run-to-run spread
    rgb8 = rt.resolve_counts(s, c).astype(np.float64)
    mu, sd = g["rgb8_mean"].astype(np.float64), g["rgb8_std"].astype(np.float64)
    z = np.abs(rgb8 - mu) / (sd * np.sqrt(1 + 1 / runs) + 1.5)
    assert (z <= 4).mean() >= 0.99, f"{name}: {(z <= 4).mean():.4f} of the bytes within 4 sigma of the reference runs"
    assert abs(rgb8.mean() - mu.mean()) <= 1.0, f"{name}: mean byte {rgb8.mean():.2f} vs reference {mu.mean():.2f}"


def test_stop_rule_against_reference_pixel_state(rt, golden_dir):
    """rtb_pixel_state_fold (the replay kernel of the adaptive renderer) on 4096 given sample sequences against the
    reference's own RecordSample / IsConverged (tests/golden/pixel_state.npz): the stop count is the reference's except
    where the decision is within float rounding of the threshold; sums agree where the counts do."""
    from test_oracle_port import _pixel_state_golden
    g, x = _pixel_state_golden(golden_dir)
    for i, (thr, mn) in enumerate(g["settings"]):
        s, q, c = rt.pixel_state_fold(x, rel_threshold=float(thr), min_samples=int(mn))
        ref_n = g[f"stop_{i}"]
        same = c.astype(np.int64) == ref_n
        assert same.mean() >= 0.995, f"setting {i}: only {same.mean():.4f} of the stop counts equal the reference's"
        ref_sum = g[f"sum_{i}"]
        assert np.all(np.abs(s[same] - ref_sum[same]) <= 2e-5 * np.maximum(np.abs(ref_sum[same]), 1e-3))
        # m2 = sumsq - sum^2 / n
        n = ref_n[same].astype(np.float64)[:, None]
        m2 = q[same].astype(np.float64) - s[same].astype(np.float64) ** 2 / n
        ref_m2 = g[f"m2_{i}"][same]
        scale = np.maximum(q[same].astype(np.float64), 1e-6)
        assert np.all(np.abs(m2 - ref_m2) <= 1e-4 * scale)
        # a different count means the rule was decided the other way at ONE sample: both stop where their own arithmetic says
        assert np.all(c >= int(mn)) and np.all(c <= x.shape[1])


def test_continue_with_keep_accum(rt, scenes):
    ds = scenes["cornell"]
    cam = rt.camera_init(rt.camera_config(**PATHB_CAMERAS["cornell"]))
    one = ds.render_adaptive(cam, max_spp=32, max_depth=20, min_samples=1000)
    a = ds.render_adaptive(cam, max_spp=12, max_depth=20, min_samples=1000)
    b = ds.render_adaptive(cam, max_spp=20, max_depth=20, min_samples=1000, spp_begin=12, state=a[:3])
    assert (b[2] == 32).all() and (one[2] == 32).all()
    assert np.abs(b[0] - one[0]).max() <= 2e-5 * one[0].max()
    assert np.abs(b[1] - one[1]).max() <= 2e-5 * one[1].max()


def test_resolve_counts_matches_write_color(rt):
    rng = np.random.default_rng(5)
    cnt = rng.integers(0, 200, size=(37, 53)).astype(np.uint32)
    s = (rng.random((37, 53, 3)) ** 3 * 2.0 * np.maximum(cnt, 1)[..., None]).astype(np.float32)
    out = rt.resolve_counts(s, cnt)
    lin = np.where(cnt[..., None] > 0, s.astype(np.float64) / np.maximum(cnt, 1).astype(np.float32).astype(np.float64)[..., None], 0.0)
    ref = (256 * np.clip(np.sqrt(np.maximum(lin, 0)), 0.0, 0.999)).astype(np.int32).astype(np.uint8)
    assert np.array_equal(out, ref)


def test_adaptive_argument_errors(rt, scenes):
    ds = scenes["cornell"]
    cam = rt.camera_init(rt.camera_config(**PATHB_CAMERAS["cornell"]))
    with pytest.raises(rt.RtbError):
        ds.render_adaptive(cam, max_spp=8, max_depth=4, flags=rt.RENDER_WAVEFRONT)
    with pytest.raises(rt.RtbError):
        ds.render_adaptive(cam, max_spp=8, max_depth=4, criterion=7)
    with pytest.raises(rt.RtbError):
        ds.render_adaptive(cam, max_spp=8, max_depth=4, min_samples=0)
