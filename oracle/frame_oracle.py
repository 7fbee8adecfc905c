"""numpy restatement of the reference's frame assembly after image decode -- TEST INFRASTRUCTURE ONLY.

Follows get_x_frame / get_rgbd_frame (ViPT/lib/train/dataset/depth_utils.py:7-132) for the two modes ViPT's
test drivers use: 'rgbrgb' (RGBT / RGBE, :128-130) and 'rgbcolormap' (RGBD, :119-123, depth clip :19-24).
Third-party arithmetic restated here (OpenCV 4.x, absent from /root/reference):
  cv2.normalize(NORM_MINMAX, 0, 255) on CV_16U : scale = 255 / (max - min), shift = -min * scale in double,
        then convertTo(CV_16U) in float: saturate_cast<ushort>(fma(src, (float)scale, (float)shift)), round half even
  cv2.applyColorMap(COLORMAP_JET)               : 256-entry table, BGR order (tests/golden/jet_lut.npy, dumped from cv2)
Pinned by tests/golden/frames_golden.npz, written by tests/golden/make_golden.py from the unmodified reference
functions (real cv2) in the build container.
"""
import os

import numpy as np

_LUT = None


def jet_lut_bgr():
    global _LUT
    if _LUT is None:
        _LUT = np.load(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden",
                                    "jet_lut.npy"))
    return _LUT


def normalize_minmax_u16(dp):
    """cv2.normalize(dp, None, 0, 255, NORM_MINMAX) for a uint16 image."""
    smin, smax = float(dp.min()), float(dp.max())
    scale = 255.0 * (1.0 / (smax - smin) if smax - smin > np.finfo(np.float64).eps else 0.0)
    shift = 0.0 - smin * scale
    a, b = np.float32(scale), np.float32(shift)
    v = (dp.astype(np.float64) * np.float64(a) + np.float64(b)).astype(np.float32)     # fused multiply-add
    return np.clip(np.rint(v), 0, 65535).astype(np.uint16)


def assemble(rgb_bgr, x, dtype="rgbcolormap", depth_clip=False):
    """rgb_bgr: uint8 (H,W,3) as cv2.imread returns it; x: uint16 (H,W) depth or uint8 (H,W,3) BGR."""
    rgb = rgb_bgr[:, :, ::-1]                                   # cv2.cvtColor(BGR2RGB), depth_utils.py:78-79
    if dtype == "rgbrgb":
        return np.ascontiguousarray(np.concatenate([rgb, x[:, :, ::-1]], axis=2))     # :128-130
    if dtype != "rgbcolormap":
        raise NotImplementedError(dtype)
    dp = x.copy()
    if depth_clip:                                              # :84-86
        max_depth = min(np.median(dp) * 3, 10000)
        dp[dp > max_depth] = max_depth
    dp8 = normalize_minmax_u16(dp).astype(np.uint8)             # :119-120
    colormap = jet_lut_bgr()[dp8]                               # :121 (BGR order, merged as is)
    return np.ascontiguousarray(np.concatenate([rgb, colormap], axis=2))             # :122
