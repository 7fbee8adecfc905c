"""vipt_op_assemble_rgbd_noalloc (no counterpart in the reference): the
'rgbcolormap' frame assembly with its histogram scratch in a static __device__ array of the library, for callers that
cannot call cudaMalloc.  Checked through the C ABI against the existing workspace entry point, bit for bit (uint8
output, the tolerance the frame-assembly tests use), on two calls in a row with different images -- the second call
would see the first call's histogram if the shared array were not cleared.  Collected last on purpose (file name)."""
import ctypes

import numpy as np
import pytest
import torch

from vipt_b200 import _lib

pytestmark = pytest.mark.gpu


def _case(seed, H, W, dev):
    rng = np.random.default_rng(seed)
    bgr = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
    yy, xx = np.mgrid[0:H, 0:W]
    depth = (700 + 13 * seed + 21 * yy + 9 * xx + rng.integers(0, 500, (H, W))).astype(np.uint16)
    depth[rng.random((H, W)) < 0.04] = 0
    depth[rng.random((H, W)) < 0.02] = 61000
    return torch.from_numpy(bgr).to(dev), torch.from_numpy(depth.view(np.int16)).to(dev)


def test_noalloc_matches_workspace_entry_point_twice():
    L = _lib.lib()
    dev = torch.device("cuda", torch.cuda.current_device())
    ws = torch.zeros((65536 + 4) * 4, dtype=torch.uint8, device=dev)
    sp = _lib.stream_ptr()
    for call, (seed, H, W, clip) in enumerate(((1, 360, 640, 1), (2, 97, 131, 0), (3, 360, 640, 1))):
        bgr, dp = _case(seed, H, W, dev)
        want = torch.empty((H, W, 6), dtype=torch.uint8, device=dev)
        got = torch.full((H, W, 6), 7, dtype=torch.uint8, device=dev)
        _lib.check(L.vipt_op_assemble_rgbd(_lib.ptr(bgr), _lib.ptr(dp), H, W, clip, _lib.ptr(want), _lib.ptr(ws), sp),
                   "vipt_op_assemble_rgbd")
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info()[0]
        _lib.check(L.vipt_op_assemble_rgbd_noalloc(_lib.ptr(bgr), _lib.ptr(dp), H, W, clip, _lib.ptr(got), sp),
                   "vipt_op_assemble_rgbd_noalloc")
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info()[0] == free0, "the no-allocation entry point changed the free device memory"
        assert torch.equal(got, want), "call %d: %d differing bytes" % (call, int((got != want).sum()))


def test_noalloc_rejects_oversized_images():
    L = _lib.lib()
    dev = torch.device("cuda", torch.cuda.current_device())
    t = torch.zeros(64, dtype=torch.uint8, device=dev)
    rc = L.vipt_op_assemble_rgbd_noalloc(_lib.ptr(t), _lib.ptr(t), 4097, 4096, 0, _lib.ptr(t), _lib.stream_ptr())
    assert rc == -1 and b"VIPT_NOALLOC_MAX_PIXELS" in L.vipt_last_error()
    torch.cuda.synchronize()
