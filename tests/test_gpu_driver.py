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


def _write_lasher(home, name, fr, gt, n_infrared=None):
    import cv2
    os.makedirs(home + "/%s/visible" % name, exist_ok=True); os.makedirs(home + "/%s/infrared" % name, exist_ok=True)
    for t in range(len(fr)):
        cv2.imwrite(home + "/%s/visible/%04d.jpg" % (name, t), fr[t][:, :, 2::-1])
        if n_infrared is None or t < n_infrared:
            cv2.imwrite(home + "/%s/infrared/%04d.jpg" % (name, t), fr[t][:, :, :2:-1])
    np.savetxt(home + "/%s/visible.txt" % name, gt, delimiter=",")


def test_driver_failure_and_length_semantics(tmp_path):
    """(a) a sequence whose first box is degenerate raises 'Too small bounding box.' in the reference and only THAT
    sequence is lost (one process per sequence, test_rgbt_mgpus.py:66-117): here it is dropped from its group, logged,
    its file is not written, the rest of the group is tracked;  (b) a sequence is as long as the shorter of its two
    image lists (`zip`, :98)."""
    pytest.importorskip("cv2")
    sd = synth.weight_set("peaked", "vipt_deep")
    home = str(tmp_path / "lasher")
    T = 4
    seqs = {}
    for k, name in enumerate(("ok0", "bad", "short")):
        fr, box0, _ = synth.synthetic_sequence(80 + k, T, "rgbt")
        gt = np.tile(np.asarray(box0, float), (T, 1))
        if name == "bad":
            gt[:, 2] = 0.0                                   # zero-width box: crop size 0
        _write_lasher(home, name, fr, gt, n_infrared=3 if name == "short" else None)
        seqs[name] = gt
    out = str(tmp_path / "results")
    res = driver.run_sequences(sorted(seqs), home, "LasHeR", sd, out, kind="rgbt", batch=3)
    assert sorted(res) == ["ok0", "short"]
    assert list(driver.run_sequences.failed) == ["bad"] and "Too small" in driver.run_sequences.failed["bad"]
    assert not os.path.exists(os.path.join(out, "bad.txt"))
    assert np.loadtxt(os.path.join(out, "ok0.txt")).shape == (T, 4)
    assert np.loadtxt(os.path.join(out, "short.txt")).shape == (3, 4)        # 4 visible, 3 infrared frames -> 3
    # the survivors equal a run without the failing sequence
    out2 = str(tmp_path / "results2")
    res2 = driver.run_sequences(["ok0", "short"], home, "LasHeR", sd, out2, kind="rgbt", batch=2)
    for n in ("ok0", "short"):
        np.testing.assert_array_equal(res[n], res2[n])


def test_rgbe_driver_visevent_layout(tmp_path):
    """RGBE driver on the GPU: VisEvent layout (vis_imgs / event_imgs *.bmp, groundtruth.txt, absent_label.txt with the
    target absent in the first frames), result file in the reference's '%.14f' comma format (test_rgbe_mgpus.py:83)."""
    cv2 = pytest.importorskip("cv2")
    sd = synth.weight_set("peaked", "vipt_deep")
    home = str(tmp_path / "visevent")
    T, first = 6, 2
    fr, box0, _ = synth.synthetic_sequence(90, T, "rgbe")
    d = home + "/seqA"
    os.makedirs(d + "/vis_imgs"); os.makedirs(d + "/event_imgs")
    for t in range(T):
        cv2.imwrite(d + "/vis_imgs/frame%04d.bmp" % t, fr[t][:, :, 2::-1])
        cv2.imwrite(d + "/event_imgs/frame%04d.bmp" % t, fr[t][:, :, :2:-1])
    np.savetxt(d + "/groundtruth.txt", np.tile(np.asarray(box0, float), (T, 1)), delimiter=",")
    np.savetxt(d + "/absent_label.txt", np.array([0] * first + [1] * (T - first)))
    out = str(tmp_path / "results")
    res = driver.run_sequences(["seqA"], home, "VisEvent", sd, out, kind="rgbe", batch=1)
    assert res["seqA"].shape == (T - first, 4)
    txt = open(os.path.join(out, "seqA.txt")).read().splitlines()
    assert len(txt) == T - first and all(len(l.split(",")) == 4 and len(l.split(",")[0].split(".")[1]) == 14 for l in txt)
    trk = BatchedViPTTracker(sd, 1, "vipt_deep", use_graph=False)
    bmp = lambda sub, t: d + "/%s/frame%04d.bmp" % (sub, t)
    trk.initialize([frames.get_x_frame(bmp("vis_imgs", first), bmp("event_imgs", first), dtype="rgbrgb")], [list(box0)])
    for t in range(first + 1, T):
        b, _ = trk.track([frames.get_x_frame(bmp("vis_imgs", t), bmp("event_imgs", t), dtype="rgbrgb")])
        np.testing.assert_allclose(res["seqA"][t - first], b[0], rtol=0, atol=1e-9)


class _StubVOT:
    """The part of ViPT/lib/test/vot/vot.py's VOT handle that the tracker loop touches, over a list of image pairs."""

    def __init__(self, pairs, region):
        self._pairs, self._region, self._i = pairs, region, 0
        self.reports = []

    def region(self):
        return self._region

    def frame(self):
        if self._i >= len(self._pairs):
            return None
        self._i += 1
        return list(self._pairs[self._i - 1])

    def report(self, region, confidence=None):
        self.reports.append((region, confidence))


def test_vot_loop_rgbd(tmp_path):
    """the VOT loop of lib/test/vot/vipt_class.py:51-98 (region -> frame -> get_rgbd_frame
    'rgbcolormap' with depth clip -> initialize / track -> report(Rectangle, score)) driven by a stub handle; boxes equal
    the batched tracker fed with the same assembled frames."""
    cv2 = pytest.importorskip("cv2")
    from vipt_b200 import vot_driver
    T = 6
    fr, box0, _ = synth.synthetic_sequence(95, T, "rgbd")
    rng = np.random.default_rng(3)
    pairs = []
    for t in range(T):
        c, dp = str(tmp_path / ("c%03d.png" % t)), str(tmp_path / ("d%03d.png" % t))
        depth = (900 + 8 * fr[t][:, :, 3].astype(np.int32) + rng.integers(0, 40, fr[t].shape[:2])).astype(np.uint16)
        cv2.imwrite(c, fr[t][:, :, 2::-1]); cv2.imwrite(dp, depth)
        pairs.append((c, dp))

    class P:
        pass
    params = P()
    params.cfg = {"TRAIN": {"PROMPT": {"TYPE": "vipt_deep"}}}
    params.state_dict = synth.weight_set("peaked", "vipt_deep")
    params.template_factor, params.template_size, params.search_factor, params.search_size = 2.0, 128, 4.0, 256
    params.save_all_boxes = False
    handle = _StubVOT(pairs, vot_driver.Rectangle(*box0))
    n = vot_driver.run_vot_exp(params, handle, out_conf=True)
    assert n == T - 1 and len(handle.reports) == T - 1
    trk = BatchedViPTTracker(params.state_dict, 1, "vipt_deep", use_graph=True)
    load = lambda p: frames.get_rgbd_frame(p[0], p[1], dtype="rgbcolormap", depth_clip=True)
    trk.initialize([load(pairs[0])], [list(box0)])
    for t in range(1, T):
        b, s = trk.track([load(pairs[t])])
        region, conf = handle.reports[t - 1]
        assert isinstance(region, vot_driver.Rectangle)
        np.testing.assert_array_equal(np.asarray(region), b[0])
        assert conf == float(s[0])
