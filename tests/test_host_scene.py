"""Host-side logic (CPU): the SAH builder reproduces the reference's trees, the flattener's
layout invariants hold, camera derivation matches the reference, loaders behave."""
import ctypes as C
import hashlib

import numpy as np
import pytest

import helpers

SCENES = {
    "cornell": lambda rt: rt.HostScene.build("cornell"),
    "rtiow": lambda rt: rt.HostScene.build("rtiow", (11, 1)),
    "bunny": lambda rt: rt.HostScene.build("bunny", (50,)),
    "earth_bubbles": lambda rt: rt.HostScene.build("earth_bubbles"),
    "mixed": lambda rt: rt.HostScene.from_records(*helpers.mixed_scene()),
}


def digest(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


@pytest.mark.parametrize("name", list(SCENES))
def test_sah_tree_identical_to_reference_golden(rt, golden_dir, name):
    """rt::geom::Bvh vs the reference's Bvh::nodes()/prim_indices() (src/geom/bvh.h:134-136), bit for bit."""
    hs = SCENES[name](rt)
    g = np.load(golden_dir / f"bvh_{name}.npz")
    assert digest(hs.export()[0]) == str(g["scene"]), "scene builder output changed; regenerate goldens"
    assert hs.bvh_count() == int(g["count"])
    for k in range(hs.bvh_count()):
        nodes, idx = hs.bvh_dump(k)
        if f"nodes{k}" in g:
            assert np.array_equal(nodes, g[f"nodes{k}"])
            assert np.array_equal(idx, g[f"idx{k}"])
        else:
            assert tuple(g[f"shape{k}"]) == nodes.shape
            assert digest(nodes) == str(g[f"sha_nodes{k}"]) and digest(idx) == str(g[f"sha_idx{k}"])


@pytest.mark.parametrize("name", list(SCENES))
def test_sah_tree_identical_to_compiled_reference(rt, name):
    from oracle import ref
    if not ref.available():
        pytest.skip("oracle/_ref not built (no /root/reference here)")
    hs = SCENES[name](rt)
    prims, mats, texs, img = hs.export()
    rs = ref.RefScene(prims, mats, texs, img)
    assert rs.bvh_count() == hs.bvh_count()
    for k in range(hs.bvh_count()):
        n1, i1 = hs.bvh_dump(k)
        n2, i2 = rs.bvh_dump(k)
        assert np.array_equal(n1, n2) and np.array_equal(i1, i2)


def test_bunny_tree_shape_matches_survey(rt):
    hs = SCENES["bunny"](rt)
    nodes, idx = hs.bvh_dump(0)  # the mesh BVH
    assert len(idx) == 69451
    leaves = nodes[nodes[:, 8] == 1]
    assert (nodes.shape[0], leaves.shape[0]) == (40539, 20270) or nodes.shape[0] in (40539, 40541)
    assert leaves[:, 7].max() <= 6  # SAH-refused leaves may exceed 4


@pytest.mark.parametrize("name", list(SCENES))
def test_flattened_layout_invariants(rt, name):
    hs = SCENES[name](rt)
    d = hs.desc
    nodes, refs = hs.nodes(), hs.prim_refs()
    assert d.n_nodes % 2 == 0 and d.n_top_nodes % 2 == 0 and d.n_top_nodes <= d.n_nodes
    internal = (nodes["b"] == 0) & (nodes["lo"][:, 0] <= nodes["hi"][:, 0])
    # children are an aligned sibling pair further down the array and inside the parent's box
    a = nodes["a"][internal]
    assert np.all(a % 2 == 0) and np.all(a > np.nonzero(internal)[0]) and np.all(a + 1 < d.n_nodes)
    for side in (0, 1):
        child = nodes[a + side]
        real = child["lo"][:, 0] <= child["hi"][:, 0]
        assert np.all(child["lo"][real] >= nodes["lo"][internal][real]) and np.all(child["hi"][real] <= nodes["hi"][internal][real])
    leaf = nodes["b"] > 0
    assert nodes["b"][leaf].max() <= 63
    # every primitive is referenced by exactly one leaf entry
    types, idx = refs >> 30, refs & 0x3FFFFFFF
    for t, n in ((0, d.n_spheres), (1, d.n_triangles), (2, d.n_rects)):
        assert np.array_equal(np.sort(idx[types == t]), np.arange(n))
    # tables are in leaf order: refs of one type appear with increasing index
    for t in (0, 1, 2):
        assert np.all(np.diff(idx[types == t]) == 1)
    gid2ref = hs.gid_to_ref()
    assert len(gid2ref) == d.n_spheres + d.n_triangles + d.n_rects and len(set(gid2ref.tolist())) == len(gid2ref)


def test_oversize_leaf_is_split_and_ranked(rt):
    R = helpers.Records()
    m = R.lambertian(0.5, 0.5, 0.5)
    R.begin(helpers.EX_BVH_BEGIN)
    for k in range(200):  # identical centroids -> the SAH builder must emit one 200-primitive leaf
        R.sphere((0, 0, 0), 1.0 + k * 1e-3, m)
    R.end(helpers.EX_BVH_END)
    hs = rt.HostScene.from_records(*R.arrays())
    nodes, idx = hs.bvh_dump(0)
    assert nodes.shape[0] == 1 and nodes[0, 7] == 200
    dn = hs.nodes()
    assert dn["b"].max() <= 63 and hs.desc.n_spheres == 200


def test_records_round_trip(rt):
    hs = SCENES["mixed"](rt)
    prims, mats, texs, img = hs.export()
    hs2 = rt.HostScene.from_records(prims, mats, texs, img)
    p2, m2, t2, _ = hs2.export()
    assert np.array_equal(prims, p2) and np.array_equal(mats, m2) and np.array_equal(texs, t2)
    assert np.array_equal(hs.nodes(), hs2.nodes()) and np.array_equal(hs.prim_refs(), hs2.prim_refs())


def test_camera_frame_matches_reference_rays(rt, golden_dir):
    """Reference Camera::GetRay samples (camera.h:134-144) must fall inside the pixel footprint the
    host mirror derives: dir = pixel00 + (i+ox) du + (j+oy) dv - origin with ox,oy in [-0.5,0.5)."""
    g = np.load(golden_dir / "camera_rays.npz")
    import ctypes as C
    for name in ("cornell", "bunny", "mixed"):  # presets without defocus: origin == centre
        cfg = g[f"cfg_{name}"]
        c = rt.camera_config(aspect_ratio=cfg[0], image_width=int(cfg[1]), vfov=cfg[4], lookfrom=cfg[5:8], lookat=cfg[8:11],
                             vup=cfg[11:14], defocus_angle=cfg[14], focus_dist=cfg[15])
        cam = rt.camera_init(c)
        p00, du, dv, cen = (np.array(v, np.float64) for v in (cam.pixel00, cam.delta_u, cam.delta_v, cam.center))
        ij, rays = g[f"ij_{name}"], g[f"rays_{name}"]
        assert np.allclose(rays[:, :3], cen, atol=1e-6)
        resid = rays[:, 3:] + rays[:, :3] - p00 - ij[:, :1] * du - ij[:, 1:] * dv
        A = np.stack([du, dv], axis=1)
        off, *_ = np.linalg.lstsq(A, resid.T, rcond=None)
        assert np.all(off >= -0.5 - 1e-4) and np.all(off < 0.5 + 1e-4)
        assert np.abs(A @ off - resid.T).max() < 1e-5 * max(1.0, np.abs(rays).max())
    cfg = g["cfg_rtiow"]
    assert int(cfg[1] / cfg[0]) == rt.camera_init(rt.camera_config(aspect_ratio=cfg[0], image_width=int(cfg[1]))).height


def test_camera_presets(rt):
    c = rt.camera_preset(rt.ASSET_DIR / "cameras.json", "cornell")
    assert (c.image_width, c.samples_per_pixel, c.max_depth, c.aspect_ratio, c.focus_dist) == (600, 100, 20, 1.5, 9.0)
    d = rt.camera_preset(rt.ASSET_DIR / "cameras.json", "default")
    assert d.defocus_angle == 0.0 and d.focus_dist == 10.0 and tuple(d.lookfrom) == (13.0, 2.0, 3.0)  # defaults, camera.h:26-37
    with pytest.raises(rt.RtbError):
        rt.camera_preset(rt.ASSET_DIR / "cameras.json", "nope")
    w = rt.camera_preset(rt.ASSET_DIR / "cameras_readme.json", "wide")
    cam = rt.camera_init(w)
    assert (cam.width, cam.height) == (800, 450)


def test_image_table_matches_reference_pipeline(rt):
    """scene::Image byte table: stbi_loadf gamma 2.2 then x255.999 (image.cc:16-73)."""
    from oracle import ref
    if not ref.available():
        pytest.skip("oracle/_ref not built")
    hs = rt.HostScene.build("earth")
    prims, mats, texs, img = hs.export()
    rs = ref.RefScene(prims, mats, texs, img)
    tex_id = int(np.nonzero(texs[:, 0] == 2)[0][0])
    rng = np.random.default_rng(0)
    uvp = np.concatenate([rng.uniform(-0.1, 1.1, (4000, 2)), np.zeros((4000, 3))], axis=1)
    want = rs.texture_value(tex_id, uvp)
    import ctypes as C
    d = hs.desc
    im = C.cast(d.images, C.POINTER(C.c_int32 * 4)).contents  # {w, h, ptr}
    w, h = im[0], im[1]
    assert (w, h) == (1024, 512)
    class Img(C.Structure):
        _fields_ = [("w", C.c_int32), ("h", C.c_int32), ("rgb", C.POINTER(C.c_uint8))]
    table = np.ctypeslib.as_array(C.cast(d.images, C.POINTER(Img)).contents.rgb, shape=(h, w, 3))
    u = np.clip(uvp[:, 0], 0, 1); v = 1 - np.clip(uvp[:, 1], 0, 1)
    i = np.minimum((u * w).astype(int), w - 1); j = np.minimum((v * h).astype(int), h - 1)
    got = table[j, i] * (1.0 / 255.0)  # texture.h:70-71 multiplies by the rounded reciprocal
    assert np.array_equal(got, want)


def test_obj_loader_counts(rt):
    hs = SCENES["bunny"](rt)
    assert hs.desc.n_triangles == 69451 and hs.desc.n_spheres == 1


@pytest.mark.parametrize("name,build", [("cornell", ("cornell", ())), ("rtiow", ("rtiow", (11, 1))), ("bunny", ("bunny", (50,))),
                                        ("earth_bubbles", ("earth_bubbles", ()))])
def test_soup_description_for_the_device_built_tree(rt, name, build):
    """RTH_BUILD_DEVICE_TREE: geom::Bvh constructors are deferred, Flattener::RunSoup emits the primitive tables in
    CREATION order with no nodes / refs / ranks; the primitives are the ones of the normally built scene."""
    soup = rt.HostScene.build(*build, device_tree=True)
    full = rt.HostScene.build(*build)
    ds, df = soup.desc, full.desc
    assert ds.flags & 1 and not df.flags
    assert not ds.nodes and not ds.prim_refs and not ds.compiled and not ds.sphere_rank and not ds.triangle_rank and not ds.rect_rank
    assert (ds.n_spheres, ds.n_triangles, ds.n_rects, ds.n_materials, ds.n_textures, ds.n_images) == \
           (df.n_spheres, df.n_triangles, df.n_rects, df.n_materials, df.n_textures, df.n_images)
    # creation order: gid -> (type, k-th primitive of that type)
    ref = soup.gid_to_ref()
    for ty in range(3):
        idx = ref[(ref >> 30) == ty] & 0x3FFFFFFF
        assert np.array_equal(idx, np.arange(idx.size))

    def table(d, ptr, n, words):
        return np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_double)), shape=(n * words,)).reshape(n, words).copy() if n else np.zeros((0, words))

    # the same primitives, the full scene's tables permuted into its leaf order: compare through the gid maps
    for (n, ptr_s, ptr_f, gid_s, gid_f, words) in ((ds.n_spheres, ds.spheres, df.spheres, ds.sphere_gid, df.sphere_gid, 4),
                                                   (ds.n_triangles, ds.triangles, df.triangles, ds.triangle_gid, df.triangle_gid, 10),
                                                   (ds.n_rects, ds.rects, df.rects, ds.rect_gid, df.rect_gid, 6)):
        if not n:
            continue
        a, b = table(ds, ptr_s, n, words), table(df, ptr_f, n, words)
        ga = np.ctypeslib.as_array(C.cast(gid_s, C.POINTER(C.c_uint32)), shape=(n,))
        gb = np.ctypeslib.as_array(C.cast(gid_f, C.POINTER(C.c_uint32)), shape=(n,))
        assert np.array_equal(a[np.argsort(ga)].view(np.uint64), b[np.argsort(gb)].view(np.uint64))
    assert soup.bvh_count() == full.bvh_count()  # the deferred Bvh objects exist, unbuilt
    for k in range(soup.bvh_count()):
        nodes, idx = soup.bvh_dump(k)
        assert nodes.shape[0] == 0
