"""RTB_SCENE_DEVICE_BUILD: the traversal tree built on the GPU (linear BVH, csrc/rtb_build.cuh) from
the primitive tables alone.  The closest hit must not depend on the tree: ids bit-exact and t bit-equal against
the reference's geom::Bvh::Hit golden vectors, renders path-for-path equal to the host-built (SAH) tree."""
import ctypes as C

import numpy as np
import pytest

import helpers

pytestmark = pytest.mark.gpu

NAMED = {"cornell": ("cornell", ()), "rtiow": ("rtiow", (11, 1)), "bunny": ("bunny", (50,)), "earth_bubbles": ("earth_bubbles", ())}


@pytest.fixture(scope="module")
def built(rt):
    out = {}
    for name, (scene, params) in NAMED.items():
        hs = rt.HostScene.build(scene, params, device_tree=True)
        out[name] = (hs, rt.DeviceScene(hs))
    return out


@pytest.mark.parametrize("name", list(NAMED))
def test_golden_hits_on_the_device_built_tree(rt, golden_dir, built, name):
    hs, ds = built[name]
    d = hs.desc
    assert d.flags & 1 and not d.nodes and not d.prim_refs and not d.triangle_rank
    g = np.load(golden_dir / f"hits_{name}.npz")
    rays = np.ascontiguousarray(g["rays"]).view(rt.RAY_DTYPE).reshape(-1)
    hits = ds.intersect(rays)
    assert helpers.check_hits(hits, g["t"], g["gid"]), "t not bit-equal to the reference"
    gid2ref = hs.gid_to_ref()
    h = g["gid"] >= 0
    assert np.array_equal(hits["ref"][h], gid2ref[g["gid"][h]])


@pytest.mark.parametrize("name", ["rtiow", "bunny", "cornell"])
def test_renders_equal_the_host_built_tree(rt, built, name):
    """Same Philox samples, same closest hits -> the same paths: equal ray counts, images equal up to the float
    order of the atomic adds."""
    scene, params = NAMED[name]
    host = rt.DeviceScene(rt.HostScene.build(scene, params))
    cam = rt.camera_init(rt.camera_config(image_width=160, vfov=60.0, lookfrom=(3, 3, 6), lookat=(0, 0.5, 0)) if name != "cornell"
                         else rt.camera_config(aspect_ratio=1.5, image_width=120, vfov=40.0, lookfrom=(5, 5, -3.6), lookat=(5, 5, 5)))
    a, sa = host.render(cam, spp=16, max_depth=10)
    b, sb = built[name][1].render(cam, spp=16, max_depth=10)
    assert sa["rays"] == sb["rays"] and sa["samples"] == sb["samples"]
    assert np.abs(a - b).max() <= 2e-5 * max(a.max(), 1.0)
    host.close()


def test_single_primitive_and_wavefront_pipeline(rt):
    hs = rt.HostScene.build("earth", device_tree=True)  # one textured sphere
    assert hs.desc.n_spheres == 1 and hs.desc.n_triangles == 0
    ds = rt.DeviceScene(hs)
    ref = rt.DeviceScene(rt.HostScene.build("earth"))
    cam = rt.camera_init(rt.camera_config(image_width=96, vfov=20.0, lookfrom=(0, 0, 12), lookat=(0, 0, 0)))
    a, sa = ref.render(cam, spp=8, max_depth=6)
    b, sb = ds.render(cam, spp=8, max_depth=6)
    c, sc = ds.render(cam, spp=8, max_depth=6, flags=rt.RENDER_WAVEFRONT, pool_size=1 << 14)
    assert sa["rays"] == sb["rays"] == sc["rays"]
    assert np.abs(a - b).max() <= 2e-5 * a.max() and np.abs(a - c).max() <= 2e-5 * a.max()


def test_large_soup_builds_and_agrees_with_host_tree(rt):
    """A 200 k-triangle + 10 k-sphere height field: device-built vs host-built tree on 1 M rays."""
    params = (317, 316, 100, 120, 2)
    dev = rt.DeviceScene(rt.HostScene.build("stress", params, device_tree=True))
    hs = rt.HostScene.build("stress", params)
    host = rt.DeviceScene(hs)
    cam = rt.camera_init(rt.camera_config(image_width=1000, vfov=60.0, lookfrom=(0, 40, 120), lookat=(0, 0, 0)))
    rays = helpers.camera_rays(rt, cam, 500_000, 5)
    h0 = host.intersect(rays)
    h1 = dev.intersect(rays)
    sec = helpers.secondary_rays(rt, rays, h0["t"], h0["prim_id"] >= 0, 6)
    assert np.array_equal(h0["prim_id"], h1["prim_id"]) and np.array_equal(h0["t"], h1["t"])
    s0, s1 = host.intersect(sec), dev.intersect(sec)
    assert np.array_equal(s0["prim_id"], s1["prim_id"]) and np.array_equal(s0["t"], s1["t"])
    assert (h0["prim_id"] >= 0).mean() > 0.3


def test_device_build_rejects_too_many_primitives(rt):
    lib = rt.load()
    hs = rt.HostScene.build("cornell", device_tree=True)
    d = rt.binding.RtbSceneDesc.from_buffer_copy(hs.desc)
    d.n_triangles = (1 << 24) + 1
    h = C.c_void_p()
    assert lib.rtb_scene_create(C.byref(d), C.byref(h)) == rt.RTB_EINVAL


def _morton_chain_scene(rt, axes, n_dup):
    """Spheres whose Morton keys have ONE bit set each (centre E * 2^-i on one axis, i = 1..21, with a far corner fixing the
    extent), plus n_dup coincident spheres at the origin: the linear BVH over them is a chain, one level per key bit
    (the depth guard must see the HEIGHT of the tree, not the length of whichever leaf chain finished last)."""
    R = helpers.Records()
    m = R.lambertian(0.5, 0.5, 0.5)
    E = float(2 ** 21)
    R.begin(helpers.EX_BVH_BEGIN)
    R.sphere((E, E, E), 0.25, m)
    for axis in axes:
        for i in range(1, 22):
            c = [0.0, 0.0, 0.0]
            c[axis] = E * 2.0 ** -i
            R.sphere(c, 0.25, m)
    for _ in range(n_dup):
        R.sphere((0.0, 0.0, 0.0), 0.25, m)
    R.end(helpers.EX_BVH_END)
    return rt.HostScene.from_records(*R.arrays())


def _as_device_build(rt, hs):
    d = rt.binding.RtbSceneDesc.from_buffer_copy(hs.desc)
    d.flags |= 1  # RTB_SCENE_DEVICE_BUILD: the tree comes from the primitive tables alone
    d.compiled, d.nodes, d.prim_refs, d.n_nodes, d.n_prim_refs, d.n_top_nodes = None, None, None, 0, 0, 0
    return d


def test_device_build_measures_the_height_of_a_degenerate_tree(rt):
    """63 single-bit keys + 64 coincident centres: a chain deeper than the 64-entry traversal stack (bvh.h:80) must be
    refused, every time (the first version reported the depth of one arbitrary leaf and let such trees through)."""
    lib = rt.load()
    hs = _morton_chain_scene(rt, (0, 1, 2), 64)
    d = _as_device_build(rt, hs)
    for _ in range(5):
        h = C.c_void_p()
        assert lib.rtb_scene_create(C.byref(d), C.byref(h)) == rt.RTB_EINVAL
        assert b"deeper than the 64-entry" in lib.rtb_last_error()


def test_deep_but_legal_device_tree_traces_like_the_host_tree(rt):
    """Two axes: a 40-odd level chain, inside the stack bound; hits equal the host-built tree's."""
    lib = rt.load()
    hs = _morton_chain_scene(rt, (0, 1), 4)
    d = _as_device_build(rt, hs)
    h = C.c_void_p()
    assert lib.rtb_scene_create(C.byref(d), C.byref(h)) == rt.RTB_OK, lib.rtb_last_error()
    host = rt.DeviceScene(hs)
    rng = np.random.default_rng(9)
    n = 20000
    rays = np.zeros(n, rt.RAY_DTYPE)
    E = float(2 ** 21)
    targets = np.concatenate([[[E, E, E], [0, 0, 0]], [[E * 2.0 ** -i, 0, 0] for i in range(1, 22)], [[0, E * 2.0 ** -i, 0] for i in range(1, 22)]])
    tgt = targets[rng.integers(0, len(targets), n)]
    rays["o"] = (tgt + rng.normal(size=(n, 3)) * 3 + np.array([0, 0, 40.0])).astype(np.float32)
    rays["d"] = (tgt + rng.normal(size=(n, 3)) * 0.2 - rays["o"]).astype(np.float32)
    rays["tmin"], rays["tmax"] = 0.001, np.inf
    a = host.intersect(rays)
    b = np.zeros(n, rt.HIT_DTYPE)
    rt.binding._check(lib, lib.rtb_intersect_batch(h, rays.ctypes.data_as(C.c_void_p), n, b.ctypes.data_as(C.c_void_p)))
    lib.rtb_scene_destroy(h)
    assert (a["prim_id"] >= 0).mean() > 0.05
    # coincident spheres tie exactly: the host tree breaks the tie by the reference's slot order, the device tree by creation order
    same = a["prim_id"] == b["prim_id"]
    tie = ~same & (a["t"] == b["t"]) & (a["prim_id"] >= 0) & (b["prim_id"] >= 0)
    assert (same | tie).all() and np.array_equal(a["t"], b["t"])


def test_large_scene_material_indices_are_validated_on_the_device(rt):
    """Above 2^18 primitives ValidateDesc's per-primitive checks run as a kernel after the upload (22 ms of host time per
    rtb_scene_create on the stress scene otherwise): a bad index must still be refused with RTB_EINVAL."""
    lib = rt.load()
    hs = rt.HostScene.build("stress", (400, 400, 60, 150, 2), device_tree=True)
    d = rt.binding.RtbSceneDesc.from_buffer_copy(hs.desc)
    assert d.n_triangles + d.n_spheres > (1 << 18)
    h = C.c_void_p()
    assert lib.rtb_scene_create(C.byref(d), C.byref(h)) == rt.RTB_OK, lib.rtb_last_error()
    lib.rtb_scene_destroy(h)
    mats = np.ctypeslib.as_array(C.cast(d.sphere_mat, C.POINTER(C.c_uint32)), shape=(d.n_spheres,)).copy()
    mats[mats.size // 2] = d.n_materials + 7
    d.sphere_mat = mats.ctypes.data
    assert lib.rtb_scene_create(C.byref(d), C.byref(h)) == rt.RTB_EINVAL
    assert b"out of range" in lib.rtb_last_error()
