"""Pins oracle/crop_oracle.py: against the committed reference outputs (cv2-backed sample_target,
ViPT/lib/train/data/processing_utils.py:14-81) and against cv2 itself where importable."""
import os

import numpy as np
import pytest

from oracle import crop_oracle as co
from vipt_b200 import synth


def test_crop_matches_reference_golden(golden_dir):
    g = np.load(os.path.join(golden_dir, "crop_golden.npz"))
    frames, _, _ = synth.synthetic_sequence(int(g["frame_seed"]), 1, "rgbt")
    im = frames[0]
    for i in range(int(g["n"])):
        box = g["box_%d" % i].tolist()
        patch, rf = co.sample_target(im, box, float(g["factor_%d" % i]), int(g["size_%d" % i]))
        assert rf == float(g["rf_%d" % i])
        assert np.array_equal(patch, g["patch_%d" % i]), "case %d not bit-exact" % i      # integer work: bit-exact
        t = co.preprocess(patch).astype(np.float64)
        assert abs(t.sum() - float(g["norm_sum_%d" % i])) <= 1e-6 * float(g["norm_abs_%d" % i])


def test_resize_bit_exact_vs_cv2():
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(3)
    for out in (128, 256):
        for s in (1, 2, 3, 31, 100, 127, 128, 129, 255, 256, 257, 300, 511, 512, 513, 777, 1024):
            src = rng.integers(0, 256, (s, s, 6), dtype=np.uint8)
            assert np.array_equal(cv2.resize(src, (out, out)), co.resize_bilinear_u8(src, out)), (s, out)


def test_geometry_edge_cases():
    # Python round is half-to-even: 0.5 -> 0, 1.5 -> 2, -0.5 -> -0
    assert co.crop_geometry([0.5 + 50, 10.0, 100.0, 100.0], 1.0, 480, 640)[0] == 50  # 50.5+50-50 -> round(50.5)=50
    assert co.crop_geometry([1.5 + 50, 10.0, 100.0, 100.0], 1.0, 480, 640)[0] == 52
    with pytest.raises(Exception, match="Too small bounding box"):
        co.crop_geometry([1.0, 1.0, 0.0, 5.0], 4.0, 480, 640)
    # the `+1` pad rule: a crop that ends exactly at the image edge still loses the last column
    im = np.full((20, 20, 6), 200, np.uint8)
    p = co.crop_padded(im, 10, 10, 10)
    assert p.shape == (10, 10, 6) and (p[:, -1] == 0).all() and (p[-1] == 0).all() and (p[:-1, :-1] == 200).all()
