"""Closest-hit parity through the C ABI (rtb_intersect_batch): primitive ids bit-exact against
geom::Bvh::Hit, t within 1e-5 relative (north_star) — in practice bit-equal, which is asserted
where the reference arithmetic is replayed exactly."""
import numpy as np
import pytest

import helpers
from test_host_scene import SCENES

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def scenes(rt):
    out = {}
    for name, make in SCENES.items():
        hs = make(rt)
        out[name] = (hs, rt.DeviceScene(hs))
    return out


@pytest.mark.parametrize("name", list(SCENES))
def test_golden_hits(rt, golden_dir, scenes, name):
    """Committed reference outputs: primary, random and bounce rays per scene."""
    g = np.load(golden_dir / f"hits_{name}.npz")
    rays = np.ascontiguousarray(g["rays"]).view(rt.RAY_DTYPE).reshape(-1)
    hits = scenes[name][1].intersect(rays)
    assert helpers.check_hits(hits, g["t"], g["gid"]), "t not bit-equal to the reference"
    gid2ref = scenes[name][0].gid_to_ref()
    h = g["gid"] >= 0
    assert np.array_equal(hits["ref"][h], gid2ref[g["gid"][h]])
    assert np.all(hits["ref"][~h] == 0xFFFFFFFF)


@pytest.mark.parametrize("name", list(SCENES))
def test_fresh_batches_against_oracle(rt, scenes, name):
    """Large seeded batches checked against the oracle run here: camera rays, random rays, two bounce generations."""
    from oracle import checker
    hs, ds = scenes[name]
    ora = checker.make_oracle(hs)
    from golden.make_golden import CAMERAS, EYES
    cam = rt.camera_init(rt.camera_config(**CAMERAS[name]))
    eye, spread, jitter = EYES[name]
    n = 200_000 if ora.kind == "reference" else 20_000
    prim = helpers.camera_rays(rt, cam, n, 21)
    batches = [prim, helpers.ray_batch(rt, n, 22, eye=eye, spread=spread, jitter=jitter)]
    t, gid = ora.intersect(prim)
    b1 = helpers.secondary_rays(rt, prim, t, gid >= 0, 23)
    batches.append(b1)
    if len(b1):
        t1, g1 = ora.intersect(b1)
        batches.append(helpers.secondary_rays(rt, b1, t1, g1 >= 0, 24))
    for rays in batches:
        if not len(rays):
            continue
        t_ref, gid_ref = ora.intersect(rays)
        assert helpers.check_hits(ds.intersect(rays), t_ref, gid_ref)


def test_adversarial_rays(rt, scenes):
    """Axis-parallel directions (1/0 in the slab test, aabb.h:98), rays aimed at shared triangle
    edges and vertices, origins on surfaces, origins inside glass, zero-length and huge directions."""
    from oracle import checker
    hs, ds = scenes["mixed"]
    ora = checker.make_oracle(hs)
    rng = np.random.default_rng(5)
    prims = hs.export()[0]
    tri = prims[prims[:, 0] == 1][:, 2:11].reshape(-1, 3, 3)
    sph = prims[prims[:, 0] == 0][:, 2:6]
    rays = []

    def add(o, d, tmin=0.001, tmax=np.inf):
        r = np.zeros(len(o), rt.RAY_DTYPE)
        r["o"], r["d"], r["tmin"], r["tmax"] = np.asarray(o, np.float32), np.asarray(d, np.float32), tmin, tmax
        rays.append(r)

    n = 20000
    o = rng.uniform(-3, 3, (n, 3)); o[:, 1] = 5
    add(o, np.tile([0, -1, 0], (n, 1)))                                  # straight down, two zero components
    add(o[:, [1, 0, 2]], np.tile([-1, 0, 0], (n, 1)))                    # along -x
    d = rng.normal(size=(n, 3)); d[:, 2] = 0
    add(rng.uniform(-3, 3, (n, 3)), d)                                   # one zero component
    k = rng.integers(0, len(tri), n)
    w = rng.uniform(0, 1, n)
    edge_pt = tri[k, 0] * w[:, None] + tri[k, 1] * (1 - w[:, None])      # points on triangle edges
    eye = np.array([0.3, 4.0, 0.2])
    add(np.tile(eye, (n, 1)), edge_pt - eye)
    add(np.tile(eye, (n, 1)), tri[k, 2] - eye)                           # aimed at vertices
    t0, g0 = ora.intersect(rays[-2])
    add(*(lambda r: (r["o"], r["d"]))(helpers.secondary_rays(rt, rays[-2], t0, g0 >= 0, 6)))  # origins on surfaces
    c = sph[rng.integers(0, len(sph), n)]
    add(c[:, :3] + rng.normal(size=(n, 3)) * 0.3 * c[:, 3:4], rng.normal(size=(n, 3)))       # origins inside spheres
    add(rng.uniform(-3, 3, (n, 3)), rng.normal(size=(n, 3)) * 1e-6)                            # tiny directions
    add(rng.uniform(-3, 3, (n, 3)), rng.normal(size=(n, 3)) * 1e6)                             # huge directions
    add(rng.uniform(-3, 3, (n, 3)), rng.normal(size=(n, 3)), tmin=0.0, tmax=2.5)               # finite interval, tmin 0
    add(np.tile([0, 8, 0], (n, 1)), edge_pt - np.array([0, 8, 0]), tmin=0.5, tmax=7.0)
    for r in rays:
        t_ref, gid_ref = ora.intersect(r)
        assert helpers.check_hits(ds.intersect(r), t_ref, gid_ref)


@pytest.mark.parametrize("name", ["mixed", "bunny"])
def test_degenerate_rays(rt, scenes, name):
    """The edges of the input domain.  Zero directions, empty / inverted intervals, a negative tmin and direction
    components that are zero, denormal or tiny next to normal ones answer like the reference.  Non-finite components
    are outside the contract (include/rtb.h: the reference itself lets NaN through its comparisons and reports hits at
    t = NaN, sphere.h:34-45): the call returns, and what it returns is a miss or a primitive of the scene."""
    from oracle import checker
    hs, ds = scenes[name]
    ora = checker.make_oracle(hs)
    rng = np.random.default_rng(77)
    n = 4096
    o = rng.uniform(-3, 3, (n, 3))
    d = rng.normal(size=(n, 3))
    pick = (np.arange(n), rng.integers(0, 3, n))
    one = d.copy(); one[pick] = rng.choice([1e-41, -1e-41, 1e-30, -3e-25, 0.0, -0.0], n)
    two = np.full((n, 3), 1e-41) * rng.choice([-1.0, 1.0], (n, 3)); two[pick] = rng.normal(size=n)

    def batch(oo, dd, tmin, tmax):
        r = np.zeros(n, rt.RAY_DTYPE)
        r["o"], r["d"], r["tmin"], r["tmax"] = oo.astype(np.float32), dd.astype(np.float32), tmin, tmax
        return r

    for r in (batch(o, np.zeros((n, 3)), 0.001, np.inf), batch(o, d, 5.0, 1.0), batch(o, d, 2.0, 2.0), batch(o, d, -10.0, np.inf),
              batch(o, one, 0.001, np.inf), batch(o, two, 0.001, np.inf)):
        t_ref, gid_ref = ora.intersect(r)
        assert helpers.check_hits(ds.intersect(r), t_ref, gid_ref)
    n_prims = int((hs.export()[0][:, 0] < 3).sum())
    bad_o, bad_d = o.copy(), d.copy()
    bad_o[pick] = rng.choice([np.nan, np.inf, -np.inf], n)
    bad_d[pick] = rng.choice([np.nan, np.inf, -np.inf], n)
    for r in (batch(bad_o, d, 0.001, np.inf), batch(o, bad_d, 0.001, np.inf), batch(o * np.nan, d * np.nan, np.nan, np.nan),
              batch(o, d, 0.001, np.nan)):
        h = ds.intersect(r)
        assert len(h) == n and np.all((h["prim_id"] >= -1) & (h["prim_id"] < n_prims))


def test_exact_ties_follow_reference_slot_order(rt, scenes):
    """Duplicated primitives: a later triangle wins an exact-t tie, an earlier sphere / rect does (bvh.h:95-106)."""
    from oracle import checker
    hs, ds = scenes["mixed"]
    ora = checker.make_oracle(hs)
    rng = np.random.default_rng(9)
    n = 30000
    targets = np.concatenate([
        np.array([0, 1, -0.3]) + rng.uniform(-0.4, 0.4, (n, 3)) * [1, 0, 1],   # duplicated triangle pair at y = 1
        np.array([0, 2.5, 0]) + rng.normal(size=(n, 3)) * 0.25,                  # duplicated spheres
        np.array([0, 3.5, 0]) + rng.uniform(-1, 1, (n, 3)) * [1, 0, 1],          # duplicated rects
    ])
    o = np.array([0.2, 2.0, 6.0]) + rng.normal(size=targets.shape) * 0.5
    rays = np.zeros(len(targets), rt.RAY_DTYPE)
    rays["o"], rays["d"], rays["tmin"], rays["tmax"] = o.astype(np.float32), (targets - o).astype(np.float32), 0.001, np.inf
    t_ref, gid_ref = ora.intersect(rays)
    hits = ds.intersect(rays)
    assert helpers.check_hits(hits, t_ref, gid_ref)
    prims = hs.export()[0]
    real = prims[prims[:, 0] < 3]
    kinds = real[gid_ref[gid_ref >= 0], 0]
    assert (kinds == 1).sum() > 1000 and (kinds == 0).sum() > 1000 and (kinds == 2).sum() > 1000


def test_empty_and_ragged_batches(rt, scenes):
    _, ds = scenes["cornell"]
    assert len(ds.intersect(np.zeros(0, rt.RAY_DTYPE))) == 0
    for n in (1, 31, 33, 1000):
        r = helpers.ray_batch(rt, n, n, eye=(5, 5, -3.6), spread=0.5, jitter=4.0)
        h = ds.intersect(r)
        assert len(h) == n and np.all((h["prim_id"] >= -1) & (h["prim_id"] < 9))


def test_miss_is_reported_like_the_reference(rt, scenes):
    _, ds = scenes["cornell"]
    r = np.zeros(64, rt.RAY_DTYPE)
    r["o"], r["d"], r["tmin"], r["tmax"] = (5, 5, -3.6), (0, 0, -1), 0.001, np.inf  # looking away from the box
    h = ds.intersect(r)
    assert np.all(h["prim_id"] == -1) and np.all(h["t"] == 0)


def test_full_size_batch_properties(rt, scenes):
    """BASELINE size (one 800x450 frame of primary rays on the bunny): size-independent properties —
    determinism, t inside the interval, shrinking tmax to the hit keeps the triangle hits."""
    hs, ds = scenes["bunny"]
    cam = rt.camera_init(rt.camera_preset(rt.ASSET_DIR / "cameras_readme.json", "wide"))
    rays = helpers.camera_rays(rt, cam, cam.width * cam.height, 31)
    h1, h2 = ds.intersect(rays), ds.intersect(rays)
    assert np.array_equal(h1, h2)
    hit = h1["prim_id"] >= 0
    assert 0.2 < hit.mean() <= 1.0 and np.all(h1["t"][hit] >= 0.001)
    tri = hit & ((h1["ref"] >> 30) == 1)
    clipped = rays.copy()
    clipped["tmax"][tri] = h1["t"][tri].astype(np.float32)  # Triangle::Hit accepts t == tmax (triangle.h:74); its t is a float
    h3 = ds.intersect(clipped)
    assert np.array_equal(h3["prim_id"][tri], h1["prim_id"][tri])


def test_scene_recreated_from_the_cached_device_image(rt, golden_dir, monkeypatch):
    """rtb_scene_create caches the scene's device image in the compiled tree (pinned): the second creation is one copy.
    Both creations, and a creation with the cache switched off, answer the golden rays identically; an image-textured
    scene keeps its per-scene texture handles out of the cache."""
    for name, build in (("bunny", ("bunny", (50,))), ("earth_bubbles", ("earth_bubbles", ()))):
        hs = rt.HostScene.build(*build)
        g = np.load(golden_dir / f"hits_{name}.npz")
        rays = np.ascontiguousarray(g["rays"]).view(rt.RAY_DTYPE).reshape(-1)
        first = rt.DeviceScene(hs)
        a = first.intersect(rays)
        second = rt.DeviceScene(hs)  # from the cached image, while the first scene is still alive
        b = second.intersect(rays)
        first.close()
        c = second.intersect(rays)
        monkeypatch.setenv("RTB_NO_ARENA_IMAGE", "1")
        plain = rt.DeviceScene(hs)
        d = plain.intersect(rays)
        monkeypatch.delenv("RTB_NO_ARENA_IMAGE")
        for h in (a, b, c, d):
            assert helpers.check_hits(h, g["t"], g["gid"])
        cam = rt.camera_init(rt.camera_config(image_width=64, vfov=30.0, lookfrom=(0, 1, 3), lookat=(0, 0, 0)))
        i1, s1 = second.render(cam, spp=8, max_depth=6)
        i2, s2 = plain.render(cam, spp=8, max_depth=6)
        assert s1["rays"] == s2["rays"] and np.abs(i1 - i2).max() <= 2e-5 * max(i1.max(), 1e-6)
