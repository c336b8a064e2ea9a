"""Image parity of the wavefront renderer (rtb_render) against the reference's Path-A estimator
(MegaKernel + DefaultSampler + Camera::GetPixel), plus renderer invariants."""
import numpy as np
import pytest

import helpers
from test_host_scene import SCENES

pytestmark = pytest.mark.gpu

GPU_SPP = 256


@pytest.fixture(scope="module")
def scenes(rt):
    out = {}
    for name, make in SCENES.items():
        hs = make(rt)
        out[name] = (hs, rt.DeviceScene(hs))
    return out


def golden_camera(rt, g):
    cfg = g["cfg"]
    c = rt.camera_config(aspect_ratio=cfg[0], image_width=int(cfg[1]), max_depth=int(cfg[3]), vfov=cfg[4], lookfrom=cfg[5:8],
                         lookat=cfg[8:11], vup=cfg[11:14], defocus_angle=cfg[14], focus_dist=cfg[15])
    return c, rt.camera_init(c)


@pytest.mark.parametrize("name", list(SCENES))
def test_image_within_3_sigma_of_reference(rt, golden_dir, scenes, name):
    """Per pixel-channel |mu_gpu - mu_ref| <= 3 sqrt(s^2/N_gpu + s^2/N_ref) for >= min(99.7 %, what the reference scores
    against itself - 0.3) of the channels, and RMSE <= 1.5 x the reference-vs-reference RMSE at equal spp."""
    g = np.load(golden_dir / f"image_{name}.npz")
    cfg, cam = golden_camera(rt, g)
    img, st = scenes[name][1].render(cam, spp=GPU_SPP, max_depth=cfg.max_depth, seed=0x3360)
    mu = img.astype(np.float64) / GPU_SPP
    ref, var, n_ref = g["mean"].astype(np.float64), g["var"].astype(np.float64), int(g["spp"])
    assert mu.shape == ref.shape
    sigma = np.sqrt(var / GPU_SPP + var / n_ref)
    # The bar is the reference's own score on the same statistic (its second render, mean_256, against the golden mean):
    # per-pixel path-tracing samples are not Gaussian and sigma is itself estimated, so the reference reaches 99.58-99.80 %
    # against itself, not 99.73 %.  The only absolute term covers float32 accumulation on zero-variance channels.
    floor = 1e-4 * np.maximum(ref, 1.0)
    ok = np.abs(mu - ref) <= 3 * sigma + floor
    ok_ref = np.abs(g["mean_256"].astype(np.float64) - ref) <= 3 * sigma + floor
    print(f"[parity] {name}: 3-sigma fraction gpu-vs-ref {ok.mean():.5f}, ref-vs-ref {ok_ref.mean():.5f}")
    assert ok.mean() >= min(0.997, ok_ref.mean() - 0.003), f"{name}: {ok.mean():.5f} of pixel-channels within 3 sigma (reference vs itself: {ok_ref.mean():.5f})"
    rmse_gpu = np.sqrt(np.mean((mu - ref) ** 2))
    rmse_ref = np.sqrt(np.mean((g["mean_256"].astype(np.float64) - ref) ** 2))
    assert rmse_gpu <= 1.5 * rmse_ref + 1e-3, f"{name}: RMSE {rmse_gpu:.5f} vs reference-vs-reference {rmse_ref:.5f}"
    # no bias: the image-wide mean agrees far tighter than any single pixel
    assert abs(mu.mean() - ref.mean()) <= 0.01 * ref.mean() + 3 * np.sqrt(np.mean(var) / (GPU_SPP * mu.size))
    # rays per camera sample agree with the reference's count
    rps = st["rays"] / st["samples"]
    assert abs(rps - float(g["rays_per_sample"])) <= 0.03 * float(g["rays_per_sample"]) + 0.02


def test_spp_split_equals_single_render(rt, scenes):
    """Multi-GPU sharding invariant on one device: rendering [0,a) and [a,n) separately and adding
    equals rendering [0,n) (Philox keys on the global sample index) up to float summation order."""
    _, ds = scenes["rtiow"]
    cam = rt.camera_init(rt.camera_config(aspect_ratio=1.777, image_width=160, vfov=45.0, lookfrom=(0, 2, 5), lookat=(0, 0, 0)))
    full, _ = ds.render(cam, spp=24, max_depth=8)
    a, _ = ds.render(cam, spp=10, max_depth=8, spp_begin=0)
    b, _ = ds.render(cam, spp=14, max_depth=8, spp_begin=10)
    assert np.allclose(a + b, full, rtol=1e-4, atol=1e-4)
    again, _ = ds.render(cam, spp=24, max_depth=8, pool_size=1 << 14)  # different pool -> different schedule, same samples
    assert np.allclose(again, full, rtol=1e-4, atol=1e-4)


def test_keep_accum_flag_accumulates(rt, scenes):
    import ctypes as C
    _, ds = scenes["cornell"]
    cam = rt.camera_init(rt.camera_config(aspect_ratio=1.5, image_width=60, vfov=40.0, lookfrom=(5, 5, -3.6), lookat=(5, 5, 5), focus_dist=9.0))
    a, _ = ds.render(cam, spp=8, max_depth=6)
    prm = rt.binding.RtbRenderParams(8, 16, 6, rt.MODE_A, 0x3360, 0, rt.RENDER_KEEP_ACCUM)
    buf = a.copy()
    st = rt.binding.RtbStats()
    rc = ds._lib.rtb_render(ds._h, C.byref(cam), C.byref(prm), buf.ctypes.data_as(C.c_void_p), C.byref(st))
    assert rc == 0
    full, _ = ds.render(cam, spp=16, max_depth=6)
    assert np.allclose(buf, full, rtol=1e-4, atol=1e-4)


def test_depth_limit_semantics(rt, scenes):
    """max_depth 0 -> black without tracing (camera.h:149-151); max_depth 1 -> only emitters and sky."""
    _, ds = scenes["cornell"]
    # vfov 80 so that the ceiling light (y = 9.99, x,z in [3,7]) is inside the frustum; the 'cornell' preset's 40 does not see it
    cam = rt.camera_init(rt.camera_config(aspect_ratio=1.5, image_width=60, vfov=80.0, lookfrom=(5, 5, -3.6), lookat=(5, 5, 5), focus_dist=9.0))
    img0, st0 = ds.render(cam, spp=4, max_depth=0)
    assert not img0.any() and st0["rays"] == 0
    img1, st1 = ds.render(cam, spp=4, max_depth=1)
    assert st1["rays"] == st1["samples"] == cam.width * cam.height * 4
    lit = img1[img1 > 0]
    assert lit.size and np.allclose(lit, np.round(lit / 15.0) * 15.0)  # only the 15,15,15 ceiling light shows


def test_depth_camera(rt, scenes):
    """DepthCamera (camera.h:215-241): first-hit distance of the normalised ray in the red channel."""
    hs, ds = scenes["cornell"]
    cfg = rt.camera_config(aspect_ratio=1.5, image_width=60, vfov=40.0, lookfrom=(5, 5, -3.6), lookat=(5, 5, 5), focus_dist=9.0)
    cam = rt.camera_init(cfg, depth_camera=True)
    img, _ = ds.render(cam, spp=16, max_depth=5)
    mu = img / 16
    assert not mu[..., 1:].any()
    centre = mu[cam.height // 2, cam.width // 2, 0]
    assert 13.0 < centre < 14.0  # back wall at z=10 seen from z=-3.6
    from oracle import checker
    ora = checker.make_oracle(hs)
    if ora.kind == "reference":
        ref_mean, _, _, _ = ora.render_a(cfg.as_oracle_cfg(), 16, 5, seed=3, want_m2=False, depth_camera=True)
        assert np.abs(mu[..., 0] - ref_mean[..., 0]).mean() < 0.05


def test_resolve_matches_write_color(rt):
    """core::write_color (color.h:18-33): sqrt gamma, clamp [0,0.999], int(256 x)."""
    rng = np.random.default_rng(2)
    s = (rng.uniform(-0.2, 1.6, (37, 53, 3)) * 10).astype(np.float32)
    s[0, 0] = (0, 10.0, 1e-30)
    got = rt.resolve(s, 10)
    lin = s.astype(np.float64) * (1.0 / 10)
    gam = np.where(lin > 0, np.sqrt(np.maximum(lin, 0)), 0.0)
    want = (256 * np.clip(gam, 0.0, 0.999)).astype(np.int64)
    assert np.array_equal(got.astype(np.int64), want)


def test_mode_b_renders_and_is_unbiased_on_diffuse_scene(rt, scenes, golden_dir):
    """Estimator B (wavefront.cc semantics: BSDF Sample + Russian roulette).  On a Lambertian-only scene
    under the sky A and B estimate the same integral up to B's sky-at-depth-limit term, so with a deep
    limit the images agree."""
    _, ds = scenes["bunny"]
    g = np.load(golden_dir / "image_bunny.npz")
    cfg, cam = golden_camera(rt, g)
    b, st = ds.render(cam, spp=GPU_SPP, max_depth=50, mode=rt.MODE_B)
    mu = b.astype(np.float64) / GPU_SPP
    a, _ = ds.render(cam, spp=GPU_SPP, max_depth=50, mode=rt.MODE_A)
    assert abs(mu.mean() - (a / GPU_SPP).mean()) < 0.01
    assert st["rays"] > st["samples"]


@pytest.mark.parametrize("name", list(SCENES))
@pytest.mark.parametrize("mode", ["A", "B"])
def test_wavefront_and_path_kernel_render_the_same_paths(rt, golden_dir, scenes, name, mode):
    """The default persistent path kernel, the pool kernel (csrc/rtb_pool.cuh) and the wavefront pipeline (RTB_RENDER_WAVEFRONT: generate /
    extend / shade+compact over queues, with and without the drain) follow the SAME paths - Philox keys
    are (pixel, sample, bounce) - so ray counts are identical and images differ only by the order of the
    float atomics."""
    g = np.load(golden_dir / f"image_{name}.npz")
    cfg, cam = golden_camera(rt, g)
    m = rt.MODE_A if mode == "A" else rt.MODE_B
    ds = scenes[name][1]
    fused, sf = ds.render(cam, spp=32, max_depth=cfg.max_depth, mode=m)
    wave, sw = ds.render(cam, spp=32, max_depth=cfg.max_depth, mode=m, flags=rt.RENDER_WAVEFRONT, pool_size=1 << 16)
    nodrain, sn = ds.render(cam, spp=32, max_depth=cfg.max_depth, mode=m, flags=rt.RENDER_WAVEFRONT | rt.RENDER_NO_DRAIN, pool_size=1 << 16)
    serial, ss = ds.render(cam, spp=32, max_depth=cfg.max_depth, mode=m, flags=rt.RENDER_SERIAL, pool_size=1 << 16)
    lane, sl = ds.render(cam, spp=32, max_depth=cfg.max_depth, mode=m, flags=rt.RENDER_POOL_PATHS)  # the pool kernel: 64 paths per warp in shared memory
    assert sl["rays"] == sf["rays"] and sl["launches"] == sf["launches"] == 1
    ordered, so = ds.render(cam, spp=32, max_depth=cfg.max_depth, mode=m, flags=rt.RENDER_COST_ORDER)  # measuring + filler + cost-ordered launch, sort kernels
    assert so["rays"] == sf["rays"] and so["launches"] == (6 if cam.width * cam.height >= 4096 else 1)  # (tiny images keep the single launch)
    assert np.abs(ordered - fused).max() <= 1e-5 * max(1.0, float(fused.max()))
    assert np.abs(lane - fused).max() <= 1e-5 * max(1.0, float(fused.max()))
    assert sf["rays"] == sw["rays"] == sn["rays"] == ss["rays"] and sf["samples"] == sw["samples"] == cam.width * cam.height * 32
    assert sf["launches"] < sw["launches"] < sn["launches"] and sw["drain_paths"] > 0 and sn["drain_paths"] == 0
    scale = max(1.0, float(fused.max()))
    for other in (wave, nodrain, serial):
        assert np.abs(other - fused).max() <= 1e-5 * scale
    assert ss["ms_extend"] > 0 and ss["ms_shade"] > 0
