"""Drop-in for the reference's frame assembly, `get_x_frame` / `get_rgbd_frame`
(ViPT/lib/train/dataset/depth_utils.py:7-132), for the modes the ViPT test drivers use:

    'rgbrgb'      RGBT / RGBE (ViPT/RGBT_workspace/test_rgbt_mgpus.py:99, RGBE_workspace/test_rgbe_mgpus.py:61)
    'rgbcolormap' RGBD        (ViPT/lib/test/evaluation/tracker.py, depthtrack / vot22 rgbd)

Image files are decoded on the host (cv2, like the reference); everything after the decode -- BGR->RGB,
depth clip at min(3 * median, 10000), min-max normalisation, JET colour map, channel merge -- runs in
libvipt_b200.so (csrc/frames.cu).  The result is a CUDA uint8 (H, W, 6) tensor that
`BatchedViPTTracker.initialize/track` accept directly (no host round trip); `.cpu().numpy()` gives the
reference's ndarray.
"""
import ctypes

import numpy as np
import torch

from . import _lib

WORKSPACE_BYTES = (65536 + 4) * 4
_ws = {}


def _workspace(device):
    key = str(device)
    if key not in _ws:
        _ws[key] = torch.empty(WORKSPACE_BYTES, dtype=torch.uint8, device=device)
    return _ws[key]


def assemble_x_frame(rgb_bgr, x, dtype="rgbcolormap", depth_clip=False, device=None):
    """rgb_bgr: uint8 (H,W,3) BGR (cv2.imread order); x: uint16 (H,W) depth ('rgbcolormap') or uint8 (H,W,3) BGR
    ('rgbrgb').  numpy arrays or tensors; returns a CUDA uint8 (H,W,6) tensor."""
    if not torch.cuda.is_available():
        raise _lib.ViptError("vipt_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
    if dtype not in ("rgbcolormap", "rgbrgb"):
        raise NotImplementedError("dtype %r (the ViPT drivers use 'rgbcolormap' and 'rgbrgb')" % (dtype,))
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    L = _lib.lib()
    a = torch.as_tensor(np.ascontiguousarray(rgb_bgr) if isinstance(rgb_bgr, np.ndarray) else rgb_bgr).to(dev).contiguous()
    if a.dtype != torch.uint8 or a.dim() != 3 or a.shape[2] != 3:
        raise ValueError("colour image must be uint8 (H, W, 3)")
    H, W = int(a.shape[0]), int(a.shape[1])
    if dtype == "rgbrgb":
        b = torch.as_tensor(np.ascontiguousarray(x) if isinstance(x, np.ndarray) else x).to(dev).contiguous()
        if b.dtype != torch.uint8 or tuple(b.shape) != (H, W, 3):
            raise ValueError("'rgbrgb' needs a second uint8 (H, W, 3) image of the same size")
    else:
        xn = x.view(np.int16) if isinstance(x, np.ndarray) and x.dtype == np.uint16 else x     # torch: reinterpret 16-bit
        b = torch.as_tensor(np.ascontiguousarray(xn) if isinstance(xn, np.ndarray) else xn).to(dev).contiguous()
        if b.element_size() != 2 or tuple(b.shape) != (H, W):
            raise ValueError("'rgbcolormap' needs a 16-bit (H, W) depth image")
    out = torch.empty((H, W, 6), dtype=torch.uint8, device=dev)
    sp = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    if dtype == "rgbrgb":
        _lib.check(L.vipt_op_assemble_rgbx(_lib.ptr(a), _lib.ptr(b), H, W, _lib.ptr(out), sp), "vipt_op_assemble_rgbx")
    else:
        _lib.check(L.vipt_op_assemble_rgbd(_lib.ptr(a), _lib.ptr(b), H, W, int(bool(depth_clip)), _lib.ptr(out),
                                           _lib.ptr(_workspace(dev)), sp), "vipt_op_assemble_rgbd")
    return out


def get_x_frame(color_path, depth_path, dtype="rgbcolormap", depth_clip=False):
    """Same signature as depth_utils.py:70; decode on the host, assembly on the device."""
    import cv2
    rgb = cv2.imread(color_path)
    dp = cv2.imread(depth_path, -1)
    if rgb is None or dp is None:
        raise FileNotFoundError("%s / %s" % (color_path, depth_path))
    return assemble_x_frame(rgb, dp, dtype, depth_clip)


get_rgbd_frame = get_x_frame      # depth_utils.py:7 -- identical for the supported modes
