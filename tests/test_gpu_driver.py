"""Batched driver end to end on a synthetic LasHeR-shaped dataset directory: result files in the reference's format,
boxes identical to single-sequence ViPTTrack runs on get_x_frame-assembled frames."""
import os

import numpy as np
import pytest
import torch

from vipt_b200 import driver, frames, synth
from vipt_b200.tracker import BatchedViPTTracker

pytestmark = pytest.mark.gpu


def test_batched_driver_writes_reference_format(tmp_path):
    cv2 = pytest.importorskip("cv2")
    sd = synth.synthetic_state_dict("vipt_deep", 0)
    home = str(tmp_path / "lasher")
    lens = {"s0": 5, "s1": 3, "s2": 4}
    gts = {}
    for k, (name, T) in enumerate(lens.items()):
        fr, box0, _ = synth.synthetic_sequence(70 + k, T, "rgbt")
        for t in range(T):
            os.makedirs(home + "/%s/visible" % name, exist_ok=True); os.makedirs(home + "/%s/infrared" % name, exist_ok=True)
            cv2.imwrite(home + "/%s/visible/%04d.jpg" % (name, t), fr[t][:, :, 2::-1])
            cv2.imwrite(home + "/%s/infrared/%04d.jpg" % (name, t), fr[t][:, :, :2:-1])
        gt = np.tile(np.asarray(box0, float), (T, 1))
        np.savetxt(home + "/%s/visible.txt" % name, gt, delimiter=",")
        gts[name] = gt
    out = str(tmp_path / "results")
    res = driver.run_sequences(sorted(lens), home, "LasHeR", sd, out, kind="rgbt", batch=2)
    assert sorted(res) == sorted(lens)
    for name, T in lens.items():
        got = np.loadtxt(os.path.join(out, name + ".txt"))
        assert got.shape == (T, 4)
        np.testing.assert_array_equal(got[0], gts[name][0])
        # the same sequence alone, frame by frame
        rgb, x, gt = driver.gen_config_rgbt(home + "/" + name, "LasHeR")
        trk = BatchedViPTTracker(sd, 1, "vipt_deep", use_graph=False)
        trk.initialize([frames.get_x_frame(rgb[0], x[0], dtype="rgbrgb")], [gt[0].tolist()])
        for t in range(1, T):
            b, _ = trk.track([frames.get_x_frame(rgb[t], x[t], dtype="rgbrgb")])
            np.testing.assert_allclose(got[t], b[0], rtol=0, atol=1e-9)
    # existing result files are skipped, like the reference (test_rgbt_mgpus.py:77-79)
    assert driver.run_sequences(sorted(lens), home, "LasHeR", sd, out, kind="rgbt", batch=2) == {}
