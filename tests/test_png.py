"""The PNG writer (host/rt/core/png.h through rth_encode_png): the file decodes, with a decoder written here from the
PNG specification (chunk CRCs checked, zlib stream inflated by Python's zlib), to the bytes that went in."""
import importlib
import struct
import zlib

import numpy as np
import pytest

binding = importlib.import_module("3360-ray-tracer_b200.binding")


def decode_png(data: bytes) -> np.ndarray:
    assert data[:8] == b"\x89PNG\r\n\x1a\n"
    at, chunks = 8, []
    while at < len(data):
        (n,) = struct.unpack(">I", data[at:at + 4])
        kind = data[at + 4:at + 8]
        body = data[at + 8:at + 8 + n]
        (crc,) = struct.unpack(">I", data[at + 8 + n:at + 12 + n])
        assert crc == zlib.crc32(kind + body) & 0xFFFFFFFF, kind
        chunks.append((kind, body))
        at += 12 + n
    assert at == len(data)
    assert [k for k, _ in chunks][0] == b"IHDR" and chunks[-1] == (b"IEND", b"")
    w, h, depth, colour, comp, flt, lace = struct.unpack(">IIBBBBB", chunks[0][1])
    assert (depth, colour, comp, flt, lace) == (8, 2, 0, 0, 0)
    raw = zlib.decompress(b"".join(b for k, b in chunks if k == b"IDAT"))  # checks the Adler-32 too
    rows = np.frombuffer(raw, dtype=np.uint8).reshape(h, 1 + 3 * w)
    assert not rows[:, 0].any()  # filter type None on every scanline
    return rows[:, 1:].reshape(h, w, 3)


@pytest.mark.parametrize("shape", [(1, 1), (2, 3), (37, 91), (450, 800), (2160, 3840)])
def test_png_round_trip(shape):
    h, w = shape
    rng = np.random.default_rng(h * 7919 + w)
    img = rng.integers(0, 256, size=(h, w, 3), dtype=np.uint8)
    data = binding.encode_png(img)
    # stored blocks: the file is the pixels + one filter byte per line + 5 bytes per 65535 + the fixed chunks
    raw = h * (3 * w + 1)
    assert len(data) == 8 + (12 + 13) + (12 + 2 + raw + 5 * -(-raw // 65535) + 4) + 12
    assert np.array_equal(decode_png(data), img)


def test_png_rejects_bad_arguments():
    with pytest.raises(ValueError):
        binding.encode_png(np.zeros((4, 4), dtype=np.uint8))
    lib = binding.load()
    import ctypes as C
    n = C.c_size_t(0)
    img = np.zeros((2, 2, 3), dtype=np.uint8)
    assert lib.rth_encode_png(img.ctypes.data, 0, 2, None, 0, C.byref(n)) != 0
    assert lib.rth_encode_png(img.ctypes.data, 2, 2, None, 0, C.byref(n)) == 0 and n.value > 0
    small = np.zeros(8, dtype=np.uint8)
    assert lib.rth_encode_png(img.ctypes.data, 2, 2, small.ctypes.data, small.size, C.byref(n)) != 0
    assert b"need" in lib.rth_last_error()
