"""host/rt/scene/jpeg_baseline.h — the product's own baseline JPEG decoder (the reference decodes textures with stb_image, which
is not part of this repository) — against the reference's texture pipeline: scene::Image -> ImageTexture::Value
(src/scene/image.cc:16-73, src/material/texture.h:59-73) compiled into oracle/_ref.  Files are made here with Pillow in every
layout the decoder claims: 4:4:4, 4:2:2, 4:2:0, greyscale, restart intervals, sizes that are not multiples of the MCU."""
import ctypes as C

import numpy as np
import pytest

import helpers

PIL = pytest.importorskip("PIL.Image")


def _picture(w, h, seed):
    rng = np.random.default_rng(seed)
    y, x = np.mgrid[0:h, 0:w]
    img = np.stack([127 + 120 * np.sin(x / 7.0 + seed), 127 + 120 * np.cos(y / 5.0), (x * 3 + y * 5) % 256], axis=-1)
    img += rng.normal(0, 12, img.shape)
    return np.clip(img, 0, 255).astype(np.uint8)


CASES = [("444", dict(subsampling=0), (64, 48)), ("422", dict(subsampling=1), (61, 35)), ("420", dict(subsampling=2), (77, 53)),
         ("420_restart", dict(subsampling=2, restart_marker_blocks=3), (50, 41)), ("grey", dict(), (33, 29)), ("tiny", dict(subsampling=2), (1, 1)),
         ("420_q30", dict(subsampling=2, quality=30), (96, 80)), ("444_optimised", dict(subsampling=0, optimize=True, quality=92), (40, 24))]


@pytest.mark.parametrize("name,opts,size", CASES)
def test_decoder_matches_the_reference_texture_pipeline(rt, tmp_path, name, opts, size):
    from oracle import ref
    if not ref.available():
        pytest.skip("oracle/_ref not built")
    w, h = size
    pic = _picture(w, h, len(name))
    im = PIL.fromarray(pic[..., 0] if name == "grey" else pic, "L" if name == "grey" else "RGB")
    path = tmp_path / f"{name}.jpg"
    im.save(path, "JPEG", **{"quality": 85, **opts})
    R = helpers.Records()
    R.texs.append([2, 0, -1, -1, 0, 0, 0, 0])  # image texture 0
    R.mats.append([0, 0, 0, 0, 0, 0, 0, 0])
    R.sphere((0, 0, 0), 1.0, 0)
    prims, mats, texs = R.arrays()
    hs = rt.HostScene.from_records(prims, mats, texs, image_paths=str(path))
    d = hs.desc

    class Img(C.Structure):
        _fields_ = [("w", C.c_int32), ("h", C.c_int32), ("rgb", C.POINTER(C.c_uint8))]
    img = C.cast(d.images, C.POINTER(Img)).contents
    assert (img.w, img.h) == (w, h), "the decoder did not load the file"
    table = np.ctypeslib.as_array(img.rgb, shape=(h, w, 3))
    # every texel through the reference: ImageTexture::Value at the texel centres
    jj, ii = np.mgrid[0:h, 0:w]
    u = (ii.ravel() + 0.5) / w
    v = 1.0 - (jj.ravel() + 0.5) / h
    rs = ref.RefScene(prims, mats, texs, str(path))
    want = rs.texture_value(0, np.stack([u, v, np.zeros_like(u), np.zeros_like(u), np.zeros_like(u)], axis=1))
    got = table.reshape(-1, 3) * (1.0 / 255.0)
    assert np.array_equal(got, want), f"{name}: {np.count_nonzero(got != want)} of {got.size} table entries differ from the reference's"


def test_unsupported_files_are_refused_with_a_message(rt, tmp_path, capfd):
    pic = _picture(40, 30, 1)
    path = tmp_path / "progressive.jpg"
    PIL.fromarray(pic, "RGB").save(path, "JPEG", progressive=True)
    R = helpers.Records()
    R.texs.append([2, 0, -1, -1, 0, 0, 0, 0])
    R.mats.append([0, 0, 0, 0, 0, 0, 0, 0])
    R.sphere((0, 0, 0), 1.0, 0)
    hs = rt.HostScene.from_records(*R.arrays(), image_paths=str(path))

    class Img(C.Structure):
        _fields_ = [("w", C.c_int32), ("h", C.c_int32), ("rgb", C.POINTER(C.c_uint8))]
    img = C.cast(hs.desc.images, C.POINTER(Img)).contents
    assert img.w == 0 and img.h == 0  # not loaded: ImageTexture::Value returns the reference's cyan (texture.h:61)
    assert "progressive JPEG is not supported" in capfd.readouterr().err
