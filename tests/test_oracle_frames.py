"""oracle/frame_oracle.py against the committed outputs of the unmodified reference get_x_frame."""
import os

import numpy as np
import pytest

from oracle import frame_oracle as fo


def test_frame_oracle_matches_reference_golden(golden_dir):
    g = np.load(os.path.join(golden_dir, "frames_golden.npz"))
    for i in range(3):
        np.testing.assert_array_equal(fo.assemble(g["bgr%d" % i], g["xbgr%d" % i], "rgbrgb"), g["rgbrgb%d" % i])
        for clip in (0, 1):
            got = fo.assemble(g["bgr%d" % i], g["depth%d" % i], "rgbcolormap", bool(clip))
            np.testing.assert_array_equal(got, g["rgbcolormap%d_clip%d" % (i, clip)])


def test_normalize_matches_cv2_when_available():
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(1)
    for _ in range(20):
        dp = rng.integers(0, 65536, (rng.integers(8, 70), rng.integers(8, 70))).astype(np.uint16)
        np.testing.assert_array_equal(fo.normalize_minmax_u16(dp), cv2.normalize(dp, None, alpha=0, beta=255,
                                                                                  norm_type=cv2.NORM_MINMAX))


def test_unsupported_dtype_raises():
    with pytest.raises(NotImplementedError):
        fo.assemble(np.zeros((4, 4, 3), np.uint8), np.zeros((4, 4), np.uint16), "rgb3x")
