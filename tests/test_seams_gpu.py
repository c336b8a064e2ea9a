"""The C++ drop-in seams used the way the reference's callers use them (tests/cpp/test_seams.cc, built by `make test_seams`):

  * integrator::B200RayIntegrator(&world).IntersectBatch(rays, hits) against the reference's CPURayIntegrator path
    (src/integrator/cpu_ray_integrator.h:18-46): EVERY HitRecord field (src/geom/hittable.h:18-34) - hit, t, p, normal,
    front_face, u, v, material present - against the unmodified reference's record for the same float ray;
  * integrator::Sampler::SamplePixel (src/integrator/sampler.h:14-20): DefaultSampler / AdaptiveSampler per pixel against
    the reference's pixel loop, and a CPU-only user sampler refused with a clear error.
"""
import os
import subprocess
from pathlib import Path

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parents[1]
EXE = ROOT / "3360-ray-tracer_b200" / "test_seams"
SCENES = {"cornell": ("cornell", ()), "rtiow": ("rtiow", (11, 1)), "bunny": ("bunny", (50,)), "earth_bubbles": ("earth_bubbles", ())}
EYES = {"cornell": ((5, 5, -3.6), 2.0, (5, 5, 5)), "rtiow": ((6, 2, 4), 3.0, (0, 0.5, 0)), "bunny": ((0, 3, 8), 2.0, (0, 0, 0)),
        "earth_bubbles": ((0, 1, 6), 2.0, (0, 0, 0))}


def _rays(name, n, seed):
    rng = np.random.default_rng(seed)
    eye, spread, target = EYES[name]
    o = (np.array(eye) + rng.normal(size=(n, 3)) * spread * 0.3).astype(np.float32)
    d = (np.array(target) + rng.normal(size=(n, 3)) * spread - o).astype(np.float32)
    return o, d


@pytest.mark.parametrize("name", list(SCENES))
def test_b200_ray_integrator_fills_the_reference_hit_record(rt, tmp_path, name):
    from oracle import checker
    if not EXE.exists():
        pytest.fail(f"{EXE} missing: run __graft_entry__.build()")
    hs = rt.HostScene.build(*SCENES[name])
    ora = checker.make_oracle(hs)
    if ora.kind != "reference":
        pytest.skip("oracle/_ref was not built")
    n = 20000
    o, d = _rays(name, n, 77)
    rays_file, out_file = tmp_path / "rays.bin", tmp_path / "hits.bin"
    np.concatenate([o, d], axis=1).astype(np.float64).tofile(rays_file)  # float32-representable doubles: the seam narrows losslessly
    r = subprocess.run([str(EXE), "hits", name, str(rays_file), str(out_file)], capture_output=True, text=True, cwd=ROOT, timeout=600,
                       env=dict(os.environ, RT_ASSETS=str(ROOT / "assets")))
    assert r.returncode == 0, r.stderr[-800:]
    got = np.fromfile(out_file, np.float64).reshape(n, 13)
    rays = np.zeros(n, rt.RAY_DTYPE)
    rays["o"], rays["d"], rays["tmin"], rays["tmax"] = o, d, 0.001, np.inf
    ref = ora.scene.intersect_records(rays)
    hit = ref[:, 0] > 0
    assert 0.2 < hit.mean() <= 1.0
    assert np.array_equal(got[:, 0], ref[:, 0]), "HitRecord::hit differs"
    g, w = got[hit], ref[hit]
    assert np.array_equal(g[:, 11], w[:, 11]), "closest primitive differs"
    assert np.array_equal(g[:, 1], w[:, 1]), "t not bit-equal to the reference's"
    # p = r.at(t) in double from the same float ray: bit-equal; the normal is the reference's formula on that p
    assert np.array_equal(g[:, 2:5], w[:, 2:5]), "p differs"
    assert np.abs(g[:, 5:8] - w[:, 5:8]).max() <= 1e-12, "normal differs"
    assert np.array_equal(g[:, 8], w[:, 8]), "front_face differs"
    assert (g[:, 12] == 1).all() and (w[:, 12] == 1).all()
    # u, v: the reference sets them for spheres (sphere.h:50) and rects (rect.h:30-31); Triangle::Hit leaves them alone
    ptype = (hs.gid_to_ref()[w[:, 11].astype(np.int64)] >> 30) & 3
    uv = ptype != rt.PRIM_TRIANGLE
    if uv.any():
        assert np.abs(g[uv, 9:11] - w[uv, 9:11]).max() <= 1e-12, "u, v differ"


def test_sampler_sample_pixel_seam(rt, tmp_path):
    from oracle import checker
    if not EXE.exists():
        pytest.fail(f"{EXE} missing: run __graft_entry__.build()")
    out = tmp_path / "sampler.txt"
    r = subprocess.run([str(EXE), "sampler", str(out)], capture_output=True, text=True, cwd=ROOT, timeout=600,
                       env=dict(os.environ, RT_ASSETS=str(ROOT / "assets")))
    assert r.returncode == 0, r.stderr[-800:]
    lines = out.read_text().strip().split("\n")
    refused = [ln for ln in lines if ln.startswith("user")]
    assert len(refused) == 1 and refused[0].startswith("user refused:") and "no CPU path" in refused[0] and "Policy()" in refused[0]
    assert any(ln.startswith("ppm P3 ") for ln in lines)
    fixed = [ln.split() for ln in lines if ln.startswith("default")]
    adapt = [ln.split() for ln in lines if ln.startswith("adaptive")]
    assert len(fixed) == 5 and len(adapt) == 5
    hs = rt.HostScene.build("cornell")
    ora = checker.make_oracle(hs)
    if ora.kind != "reference":
        pytest.skip("oracle/_ref was not built")
    cfg = rt.camera_config(aspect_ratio=1.5, image_width=120, samples_per_pixel=64, max_depth=12, vfov=40.0, lookfrom=(5, 5, -3.6),
                           lookat=(5, 5, 5), focus_dist=9.0)
    n_ref = 4096
    mean, m2, _, _ = ora.render_a(cfg.as_oracle_cfg(), n_ref, 12, seed=5, threads=len(os.sched_getaffinity(0)))
    var = m2 / (n_ref - 1)
    for rows, spp_of in ((fixed, lambda n: n), (adapt, lambda n: n)):
        for _, i, j, n, r_, g_, b_ in rows:
            i, j, n = int(i), int(j), int(n)
            got = np.array([float(r_), float(g_), float(b_)])
            sigma = np.sqrt(var[j, i] / n + var[j, i] / n_ref)
            assert np.all(np.abs(got - mean[j, i]) <= 4.5 * sigma + 1e-3), f"pixel ({i},{j}) {got} vs reference {mean[j, i]} +- {sigma}"
    assert all(int(row[3]) == 1024 for row in fixed)
    assert all(16 <= int(row[3]) <= 512 for row in adapt)
