"""pb2_trace_closest_device_noalloc (include/pine_b200.h): the closest-hit launch with its scratch in a static
__device__ array instead of a cudaMalloc'ed buffer. No counterpart in the reference; added
for callers that cannot allocate (graph capture, serving loops).

Checked through the C ABI against the existing entry point pb2_trace_closest_device on the same device rays:
geom, prim and the bits of t are IDENTICAL (the bar the closest-hit tests use), on calls made back to back on the
shared scratch with different ray counts, and against the reference's golden results. The file is named so that
pytest collects it last.
"""
import numpy as np
import pytest

import common
import pine_b200 as pb
from common import u32
from pine_b200 import _abi as A

pytestmark = pytest.mark.gpu


def _device_rays(torch, rays):
    return torch.from_numpy(np.ascontiguousarray(rays, np.float32).reshape(-1, 8)).cuda()


def _trace(torch, B, fn, rays_dev, accel=0):
    """One call of `fn` (either entry point) on torch's current stream; the caller owns every buffer."""
    n = rays_dev.shape[0]
    geom = torch.full((max(n, 1),), -7, dtype=torch.int32, device="cuda")
    prim = torch.full((max(n, 1),), -7, dtype=torch.int32, device="cuda")
    t = torch.full((max(n, 1),), -7.0, dtype=torch.float32, device="cuda")
    st = torch.cuda.current_stream().cuda_stream  # 0 (the legacy default stream) selects the context's own stream
    torch.cuda.synchronize()  # inputs and the -7 fills are in place whichever stream the launch uses
    A.check(fn(B.h, rays_dev.data_ptr(), n, accel, geom.data_ptr(), prim.data_ptr(), t.data_ptr(), st))
    torch.cuda.synchronize()
    return geom[:n].cpu().numpy(), prim[:n].cpu().numpy(), t[:n].cpu().numpy()


def _same(a, b):
    return np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and np.array_equal(u32(a[2]), u32(b[2]))


@pytest.fixture(scope="module")
def torch_cuda():
    import torch
    assert torch.cuda.is_available()
    torch.cuda.set_device(0)
    return torch


@pytest.mark.parametrize("name", ["c1_64", "mesh_48"])
def test_noalloc_equals_the_allocating_entry_point(golden, torch_cuda, name):
    torch = torch_cuda
    L = A.lib()
    g = golden(name)
    S = pb.Scene()
    common.build(S, name)
    B = pb.Backend(0)
    try:
        B.upload_scene(S)
        A.check(L.pb2_trace_closest_device_noalloc(B.h, None, 0, 0, None, None, None, None))  # set-up only
        primary, secondary = _device_rays(torch, g["primary0"]), _device_rays(torch, g["secondary"])
        # a batch large enough for many claims of the persistent warps and a long fallback list
        big = _device_rays(torch, np.tile(np.concatenate([g["primary0"], g["secondary"]]), (24, 1)))
        assert big.shape[0] <= A.NOALLOC_MAX_RAYS
        want = {k: _trace(torch, B, L.pb2_trace_closest_device, r) for k, r in
                (("primary", primary), ("secondary", secondary), ("big", big))}
        # against the reference's own results (golden vectors of the unmodified reference)
        for k, pre in (("primary", "p_"), ("secondary", "s_")):
            assert np.array_equal(want[k][0], g[pre + "geom"]) and np.array_equal(u32(want[k][2]), u32(g[pre + "t"]))
        # calls in a row on the shared static scratch: sizes go up and down, so stale list contents and counters of
        # the previous call would show
        free0 = torch.cuda.mem_get_info()[0]
        for k, r in (("big", big), ("primary", primary), ("secondary", secondary), ("big", big), ("primary", primary)):
            got = _trace(torch, B, L.pb2_trace_closest_device_noalloc, r)
            assert _same(got, want[k]), k
        # the reference-order walk (accel = 1) goes through the same entry point without scratch
        assert _same(_trace(torch, B, L.pb2_trace_closest_device_noalloc, secondary, accel=1), want["secondary"])
        # nothing was allocated by the calls themselves (outputs above come from torch's caching allocator, which
        # reuses the blocks of the first round); device free memory is what it was
        assert torch.cuda.mem_get_info()[0] >= free0 - (2 << 20)
    finally:
        B.close()


def test_noalloc_rejects_sizes_beyond_the_static_scratch(torch_cuda):
    torch = torch_cuda
    L = A.lib()
    S = pb.Scene()
    common.build(S, "cones_48")
    B = pb.Backend(0)
    try:
        B.upload_scene(S)
        n = A.NOALLOC_MAX_RAYS + 1
        dummy = torch.zeros(8, dtype=torch.float32, device="cuda")
        rc = L.pb2_trace_closest_device_noalloc(B.h, dummy.data_ptr(), n, 0, dummy.data_ptr(), dummy.data_ptr(),
                                                dummy.data_ptr(), None)
        assert rc == -1  # PB2_E_INVALID
        assert b"PB2_NOALLOC_MAX_RAYS" in L.pb2_last_error()
        # exactly the maximum is accepted
        rays = np.tile(common.ray_fan((0.0, 0.5, 0.0), 1 << 16, 3), (16, 1))
        assert len(rays) == A.NOALLOC_MAX_RAYS
        r = _device_rays(torch, rays)
        a = _trace(torch, B, L.pb2_trace_closest_device_noalloc, r)
        b = _trace(torch, B, L.pb2_trace_closest_device, r)
        assert _same(a, b)
    finally:
        B.close()


def test_noalloc_needs_a_scene_and_a_device():
    L = A.lib()
    B = pb.Backend(0)
    try:
        assert L.pb2_trace_closest_device_noalloc(B.h, None, 0, 0, None, None, None, None) == -4  # PB2_E_STATE
    finally:
        B.close()
    assert L.pb2_trace_closest_device_noalloc(None, None, 0, 0, None, None, None, None) == -1
