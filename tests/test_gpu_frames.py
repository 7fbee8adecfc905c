"""csrc/frames.cu through the C ABI against the oracle and the committed reference outputs (bit-exact)."""
import os

import numpy as np
import pytest
import torch

from oracle import frame_oracle as fo
from vipt_b200 import frames

pytestmark = pytest.mark.gpu


def test_assemble_matches_reference_golden(golden_dir):
    g = np.load(os.path.join(golden_dir, "frames_golden.npz"))
    for i in range(3):
        out = frames.assemble_x_frame(g["bgr%d" % i], g["xbgr%d" % i], "rgbrgb")
        assert out.is_cuda and out.dtype == torch.uint8
        np.testing.assert_array_equal(out.cpu().numpy(), g["rgbrgb%d" % i])
        for clip in (0, 1):
            out = frames.assemble_x_frame(g["bgr%d" % i], g["depth%d" % i], "rgbcolormap", bool(clip))
            np.testing.assert_array_equal(out.cpu().numpy(), g["rgbcolormap%d_clip%d" % (i, clip)])


@pytest.mark.parametrize("H,W", [(360, 640), (480, 630), (1, 1), (7, 3)])
def test_assemble_vs_oracle_full_size(H, W):
    rng = np.random.default_rng(H * 1000 + W)
    bgr = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
    depth = rng.integers(0, 20000, (H, W)).astype(np.uint16)
    for clip in (False, True):
        got = frames.assemble_x_frame(bgr, depth, "rgbcolormap", clip).cpu().numpy()
        np.testing.assert_array_equal(got, fo.assemble(bgr, depth, "rgbcolormap", clip))
    xb = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
    np.testing.assert_array_equal(frames.assemble_x_frame(bgr, xb, "rgbrgb").cpu().numpy(), fo.assemble(bgr, xb, "rgbrgb"))


def test_bad_args_raise():
    with pytest.raises(NotImplementedError):
        frames.assemble_x_frame(np.zeros((4, 4, 3), np.uint8), np.zeros((4, 4), np.uint16), "rgb3x")
    with pytest.raises(ValueError):
        frames.assemble_x_frame(np.zeros((4, 4, 3), np.uint8), np.zeros((5, 4), np.uint16), "rgbcolormap")
