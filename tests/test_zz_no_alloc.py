"""pt_trace_closest_device_noalloc (include/pt_b200.h): the closest-hit batch with its wave scratch in a static device
array of the library instead of cudaMalloc'ed memory.  Checked through the C ABI against the existing entry point
pt_trace_closest_device on the same device buffers, bit for bit (the bar of the closest-hit parity tests), on calls in
a row that share the array: a full-size batch on one scene, then a smaller batch on another scene whose records land
on top of the first call's.  Collected last on purpose (the file name): a failure here hides no other test under -x.
"""
import os

import numpy as np
import pytest

import isf_py
import rays as R
from conftest import GOLDEN

import path_tracer_b200 as pt

pytestmark = pytest.mark.gpu

MAX_RAYS = pt.api.TRACE_NOALLOC_MAX_RAYS


def _scene(name):
    path = os.path.join(GOLDEN, "scenes", name, "scene.isf")
    return isf_py.load_isf(path), pt.Scene.load(path)


def _both(torch, gsc, rays_np):
    """(hits of pt_trace_closest_device, hits of pt_trace_closest_device_noalloc) for one batch of host rays."""
    n = len(rays_np)
    d_rays = torch.from_numpy(np.ascontiguousarray(rays_np, np.float32)).cuda()
    outs = []
    for fn in (gsc.trace_closest_device, gsc.trace_closest_device_noalloc):
        d_out = torch.full((n * pt.abi.HIT_DTYPE.itemsize,), 0xA5, dtype=torch.uint8, device="cuda")  # poisoned
        fn(d_rays.data_ptr(), n, d_out.data_ptr())  # default stream
        torch.cuda.synchronize()
        outs.append(d_out.cpu().numpy().view(pt.abi.HIT_DTYPE).copy())
    return outs


def _assert_identical(a, b):
    assert a.tobytes() == b.tobytes(), R.compare_hits(b, a)


def test_noalloc_matches_device_entry_point_two_calls_in_a_row():
    torch = pytest.importorskip("torch")
    sa1, sc1 = _scene("head")
    sa2, sc2 = _scene("alpha_transparency")
    # call 1: the largest batch the static array takes
    r1 = R.random_rays(sa1, MAX_RAYS)
    ref1, got1 = _both(torch, sc1, r1)
    assert (ref1["prim_id"] >= 0).sum() > 1000
    _assert_identical(ref1, got1)
    assert ref1.tobytes() == sc1.trace_closest(r1).tobytes()  # and the host-pointer entry point
    # call 2, straight after: another scene, a ragged smaller batch over the first call's leftovers
    r2 = R.random_rays(sa2, 12345)
    ref2, got2 = _both(torch, sc2, r2)
    assert (ref2["prim_id"] >= 0).sum() > 100
    _assert_identical(ref2, got2)
    # call 3: the first batch again gives the first answer again, and holds no new device memory
    d_rays = torch.from_numpy(r1).cuda()
    d_out = torch.zeros(MAX_RAYS * pt.abi.HIT_DTYPE.itemsize, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    sc1.trace_closest_device_noalloc(d_rays.data_ptr(), MAX_RAYS, d_out.data_ptr())
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] >= free_before
    _assert_identical(ref1, d_out.cpu().numpy().view(pt.abi.HIT_DTYPE))


def test_noalloc_on_a_side_stream_and_single_ray():
    torch = pytest.importorskip("torch")
    sa, sc = _scene("spheres")
    rr = R.random_rays(sa, 4097)
    ref = sc.trace_closest(rr)
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        d_rays = torch.from_numpy(rr).cuda()
        d_out = torch.zeros(len(rr) * pt.abi.HIT_DTYPE.itemsize, dtype=torch.uint8, device="cuda")
        sc.trace_closest_device_noalloc(d_rays.data_ptr(), len(rr), d_out.data_ptr(), stream=st.cuda_stream)
        # one ray, same stream (ordered after the batch): the result lands in a second buffer
        d_one = torch.zeros(pt.abi.HIT_DTYPE.itemsize, dtype=torch.uint8, device="cuda")
        sc.trace_closest_device_noalloc(d_rays.data_ptr(), 1, d_one.data_ptr(), stream=st.cuda_stream)
    st.synchronize()
    _assert_identical(ref, d_out.cpu().numpy().view(pt.abi.HIT_DTYPE))
    _assert_identical(ref[:1], d_one.cpu().numpy().view(pt.abi.HIT_DTYPE))


def test_noalloc_rejects_oversize_and_bad_arguments():
    torch = pytest.importorskip("torch")
    _, sc = _scene("cube")
    d = torch.zeros(64, dtype=torch.uint8, device="cuda")
    with pytest.raises(pt.PtError) as e:
        sc.trace_closest_device_noalloc(d.data_ptr(), MAX_RAYS + 1, d.data_ptr())
    assert e.value.code == -1 and "PT_TRACE_NOALLOC_MAX_RAYS" in str(e.value) and str(MAX_RAYS) in str(e.value)
    with pytest.raises(pt.PtError):
        sc.trace_closest_device_noalloc(None, 16, d.data_ptr())
    sc.trace_closest_device_noalloc(None, 0, None)  # empty batch: nothing to do, no error
