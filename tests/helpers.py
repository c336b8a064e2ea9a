"""Shared test helpers: scene records built in Python, seeded ray batches."""
from __future__ import annotations

import numpy as np

EX_SPHERE, EX_TRIANGLE, EX_RECT, EX_BVH_BEGIN, EX_BVH_END, EX_LIST_BEGIN, EX_LIST_END = 0, 1, 2, 10, 11, 12, 13


class Records:
    """Builds the export stream (16 doubles / primitive record) the host API and the oracle read."""

    def __init__(self):
        self.prims, self.mats, self.texs = [], [], []
        self.begin(EX_LIST_BEGIN)

    def _rec(self, kind, mat=0, params=()):
        r = np.zeros(16)
        r[0], r[1] = kind, mat
        r[2:2 + len(params)] = params
        self.prims.append(r)

    def begin(self, kind):
        self._rec(kind)

    def end(self, kind):
        self._rec(kind)

    def solid(self, r, g, b):
        self.texs.append([0, -1, -1, -1, r, g, b, 0])
        return len(self.texs) - 1

    def checker(self, scale, even, odd):
        self.texs.append([1, -1, even, odd, 0, 0, 0, scale])
        return len(self.texs) - 1

    def lambertian(self, r, g, b):
        self.mats.append([0, self.solid(r, g, b), 0, 0, 0, 0, 0, 0])
        return len(self.mats) - 1

    def lambertian_tex(self, tex):
        self.mats.append([0, tex, 0, 0, 0, 0, 0, 0])
        return len(self.mats) - 1

    def metal(self, r, g, b, fuzz):
        self.mats.append([1, -1, r, g, b, fuzz, 0, 0])
        return len(self.mats) - 1

    def dielectric(self, ior):
        self.mats.append([2, -1, 0, 0, 0, 0, ior, 0])
        return len(self.mats) - 1

    def light(self, r, g, b):
        self.mats.append([3, self.solid(r, g, b), 0, 0, 0, 0, 0, 0])
        return len(self.mats) - 1

    def sphere(self, c, r, mat):
        self._rec(EX_SPHERE, mat, [*c, r])

    def triangle(self, a, b, c, mat):
        self._rec(EX_TRIANGLE, mat, [*a, *b, *c])

    def rect(self, axis, k, a0, a1, b0, b1, mat):
        self._rec(EX_RECT, mat, [axis, k, a0, a1, b0, b1])

    def arrays(self):
        prims = self.prims + [np.array([EX_LIST_END] + [0] * 15, dtype=float)]
        return (np.array(prims, dtype=np.float64).reshape(-1, 16), np.array(self.mats, dtype=np.float64).reshape(-1, 8),
                np.array(self.texs, dtype=np.float64).reshape(-1, 8))


def mixed_scene(seed=3, n_tri=300, n_sph=60):
    """Spheres + triangles + rects in nested BVHs and a bare list, with exact duplicates (t ties)."""
    rng = np.random.default_rng(seed)
    R = Records()
    m_d = R.lambertian(0.7, 0.3, 0.3)
    m_m = R.metal(0.8, 0.8, 0.9, 0.1)
    m_g = R.dielectric(1.5)
    m_l = R.light(4, 4, 4)
    m_c = R.lambertian_tex(R.checker(0.5, R.solid(0.1, 0.1, 0.1), R.solid(0.9, 0.9, 0.9)))
    # a mesh BVH with shared edges (grid of quads -> 2 triangles each), float-representable vertices
    R.begin(EX_BVH_BEGIN)
    R.begin(EX_BVH_BEGIN)
    g = int(np.sqrt(n_tri / 2))
    zs = rng.uniform(-0.3, 0.3, (g + 1, g + 1)).astype(np.float32)
    for i in range(g):
        for j in range(g):
            p = lambda a, b: (np.float32(-2 + 4 * a / g), np.float32(zs[a, b]), np.float32(-2 + 4 * b / g))  # noqa: E731
            R.triangle(p(i, j), p(i, j + 1), p(i + 1, j + 1), m_d)
            R.triangle(p(i, j), p(i + 1, j + 1), p(i + 1, j), m_m if (i + j) % 7 == 0 else m_d)
    # duplicated triangles (exact ties, later slot wins for triangles)
    R.triangle((-1, 1, -1), (1, 1, -1), (0, 1, 1), m_d)
    R.triangle((-1, 1, -1), (1, 1, -1), (0, 1, 1), m_m)
    R.end(EX_BVH_END)
    for k in range(n_sph):
        c = rng.uniform(-2.5, 2.5, 3)
        c[1] = abs(c[1]) + 0.6
        R.sphere(c, float(rng.uniform(0.1, 0.45)), [m_d, m_m, m_g][k % 3])
    # duplicated spheres (exact ties, earlier slot wins) and concentric hollow glass
    R.sphere((0, 2.5, 0), 0.5, m_g)
    R.sphere((0, 2.5, 0), 0.5, m_d)
    R.sphere((1.5, 2.0, 0.5), 0.4, m_g)
    R.sphere((1.5, 2.0, 0.5), -0.35 + 0.7, m_g)
    R.rect(1, 3.5, -1, 1, -1, 1, m_l)      # xz light
    R.rect(1, 3.5, -1, 1, -1, 1, m_d)      # duplicate rect
    R.rect(2, -3.1, -3, 3, -1, 4, m_c)     # xy wall (off the checker lattice: floor(z/0.5) must not sit on a cell boundary)
    R.rect(0, -3.0, -1, 4, -3, 3, m_d)     # yz wall
    R.end(EX_BVH_END)
    # objects straight in the world list, next to the BVH
    R.sphere((0, -1000.5, 0), 1000, m_c)
    R.triangle((3, 0, 3), (3, 2, 3), (3, 0, -3), m_d)
    return R.arrays()


def ray_batch(rt, n, seed, eye=(0.0, 2.0, 8.0), spread=3.0, jitter=2.0, tmin=0.001):
    rng = np.random.default_rng(seed)
    r = np.zeros(n, rt.RAY_DTYPE)
    r["o"] = (rng.normal(size=(n, 3)) * spread + np.array(eye)).astype(np.float32)
    r["d"] = (-r["o"] + rng.normal(size=(n, 3)) * jitter).astype(np.float32)
    r["tmin"] = tmin
    r["tmax"] = np.inf
    return r


def camera_rays(rt, cam, n, seed):
    """Primary rays the way generate builds them, on the host (pixel centre + jitter, no defocus)."""
    rng = np.random.default_rng(seed)
    i = rng.integers(0, cam.width, n)
    j = rng.integers(0, cam.height, n)
    fi = (i + rng.uniform(-0.5, 0.5, n)).astype(np.float32)
    fj = (j + rng.uniform(-0.5, 0.5, n)).astype(np.float32)
    p00, du, dv, c = (np.array(v, np.float32) for v in (cam.pixel00, cam.delta_u, cam.delta_v, cam.center))
    r = np.zeros(n, rt.RAY_DTYPE)
    r["o"] = c
    r["d"] = p00 + fi[:, None] * du + fj[:, None] * dv - c
    r["tmin"] = 0.001
    r["tmax"] = np.inf
    return r


def secondary_rays(rt, rays, t, hit, seed):
    """Bounce rays leaving the hit points of `rays` in random directions (incoherent batch)."""
    rng = np.random.default_rng(seed)
    src = rays[hit]
    p = (src["o"].astype(np.float64) + t[hit][:, None] * src["d"].astype(np.float64)).astype(np.float32)
    d = rng.normal(size=p.shape).astype(np.float32)
    r = np.zeros(len(p), rt.RAY_DTYPE)
    r["o"], r["d"], r["tmin"], r["tmax"] = p, d, 0.001, np.inf
    return r


def check_hits(hits, t_ref, gid_ref, rtol=1e-5):
    """The north-star criterion: ids bit-exact, t within 1e-5 relative.  Returns t-bit-equality too."""
    assert np.array_equal(hits["prim_id"], gid_ref), f"{np.count_nonzero(hits['prim_id'] != gid_ref)} closest-hit ids differ"
    h = gid_ref >= 0
    assert np.all(np.abs(hits["t"][h] - t_ref[h]) <= rtol * np.abs(t_ref[h]))
    return bool(np.array_equal(hits["t"][h], t_ref[h]))
