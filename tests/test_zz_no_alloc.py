"""rtb_intersect_batch_static (include/rtb.h): seam 2 for callers that must not allocate - the kernels' scratch is a file-scope
device array, the call performs no cudaMalloc / cudaFree.  No counterpart in the reference;
its answers are those of rtb_intersect_batch, which tests/test_hits_gpu.py pins to the reference's geom::Bvh::Hit.

Collected last on purpose.  Passes alone in a fresh process and after the other tests: it builds
its own scenes and only relies on the library's per-thread stream."""
import re
from pathlib import Path

import numpy as np
import pytest

import helpers

ROOT = Path(__file__).resolve().parents[1]


def test_entry_point_body_calls_no_allocator():
    src = (ROOT / "3360-ray-tracer_b200" / "csrc" / "rtb_api.cu").read_text()
    start = src.index("int rtb_intersect_batch_static(")
    body = src[start:src.index("\nint rtb_intersect_batch(", start)]
    assert "cudaGetSymbolAddress(&p, g_static_batch)" in src and "StaticBatchScratch(s->device, &scratch)" in body
    assert not re.search(r"cudaMalloc|cudaFree|cudaHostAlloc|cudaMallocHost|\bnew\b|std::vector|malloc\(", body)
    hdr = (ROOT / "include" / "rtb.h").read_text()
    assert "RTB_STATIC_BATCH_MAX 65536u" in hdr and "rtb_intersect_batch_static" in hdr


@pytest.mark.gpu
@pytest.mark.parametrize("name,params", [("bunny", (50,)), ("cornell", ())])
def test_static_entry_matches_the_allocating_one_twice_in_a_row(rt, name, params):
    ds = rt.DeviceScene(rt.HostScene.build(name, params))
    eye = (0.0, 2.0, 6.0) if name == "bunny" else (5.0, 5.0, 2.0)
    first = helpers.ray_batch(rt, rt.STATIC_BATCH_MAX, 201, eye=eye, spread=3.0, jitter=2.5)   # the full array
    second = helpers.ray_batch(rt, 20_011, 202, eye=eye, spread=2.0, jitter=4.0)              # fewer rays: the array still holds call 1
    for rays in (first, second):
        want = ds.intersect(rays)          # rtb_intersect_batch: cudaMalloc'ed staging, the scene's own cursor
        got = ds.intersect_static(rays)    # the same kernels on the static scratch
        # the tolerance of the closest-hit tests: ids bit-exact, t bit-equal (tests/helpers.check_hits)
        assert np.array_equal(got["prim_id"], want["prim_id"])
        assert np.array_equal(got["ref"], want["ref"])
        assert np.array_equal(got["t"], want["t"])
        assert 0.05 < (got["prim_id"] >= 0).mean()
    # the same batch again: the cursor of the previous call was cleared, nothing else carried over
    again = ds.intersect_static(second)
    assert np.array_equal(again, ds.intersect(second))
    ds.close()


@pytest.mark.gpu
def test_static_entry_rejects_batches_beyond_its_maximum(rt):
    ds = rt.DeviceScene(rt.HostScene.build("cornell"))
    rays = helpers.ray_batch(rt, rt.STATIC_BATCH_MAX + 1, 203, eye=(5.0, 5.0, 2.0))
    with pytest.raises(rt.RtbError) as e:
        ds.intersect_static(rays)
    assert e.value.code == rt.RTB_EINVAL and "RTB_STATIC_BATCH_MAX" in str(e.value) and "65536" in str(e.value)
    assert len(ds.intersect_static(rays[:0])) == 0  # an empty batch is a no-op
    hits = ds.intersect_static(rays[:1000])           # and the refusal left the entry point usable
    assert np.array_equal(hits, ds.intersect(rays[:1000]))
    ds.close()
