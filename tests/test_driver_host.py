"""Host logic of the batched benchmark drivers (no GPU): dataset layouts and byte-compatible result files
(ViPT/RGBT_workspace/test_rgbt_mgpus.py:20-117, ViPT/RGBE_workspace/test_rgbe_mgpus.py:18-84)."""
import os

import numpy as np
import pytest

from vipt_b200 import driver


def _touch(path):
    os.makedirs(os.path.dirname(path), exist_ok=True)
    open(path, "wb").close()


def test_gen_config_layouts(tmp_path):
    s = str(tmp_path / "lasher" / "seqA")
    for i in (2, 0, 1):
        _touch(s + "/visible/v%04d.jpg" % i); _touch(s + "/infrared/i%04d.jpg" % i)
    _touch(s + "/visible/notes.txt")
    np.savetxt(s + "/visible.txt", np.array([[1, 2, 30, 40], [2, 3, 30, 40], [3, 4, 31, 41]], float), delimiter=",", fmt="%d")
    rgb, x, gt = driver.gen_config_rgbt(s, "LasHeR")
    assert [os.path.basename(p) for p in rgb] == ["v0000.jpg", "v0001.jpg", "v0002.jpg"] and len(x) == 3
    assert rgb[0] == s + "/visible/v0000.jpg" and gt.shape == (3, 4) and gt[2, 3] == 41
    g = str(tmp_path / "gtot" / "seqB")
    for i in range(2):
        _touch(g + "/v/%d.png" % i); _touch(g + "/i/%d.png" % i)
    np.savetxt(g + "/groundTruth_v.txt", np.array([[50, 60, 10, 20], [11, 21, 51, 61]], float), delimiter=" ")
    _, _, gt = driver.gen_config_rgbt(g, "GTOT")
    np.testing.assert_array_equal(gt, [[10, 20, 40, 40], [11, 21, 40, 40]])       # corners -> xywh
    e = str(tmp_path / "visevent" / "seqC")
    for i in range(4):
        _touch(e + "/vis_imgs/%03d.bmp" % i); _touch(e + "/event_imgs/%03d.bmp" % i)
    np.savetxt(e + "/groundtruth.txt", np.arange(16, dtype=float).reshape(4, 4), delimiter=",")
    np.savetxt(e + "/absent_label.txt", np.array([0, 0, 1, 1.0]))
    rgb, ev, gt, absent = driver.gen_config_rgbe(e, "VisEvent")
    assert len(rgb) == 2 and rgb[0].endswith("002.bmp") and gt[0, 0] == 8           # trimmed to first present frame
    with pytest.raises(ValueError):
        driver.gen_config_rgbt(s, "nope")


def test_result_files_are_byte_compatible(tmp_path):
    gt = np.array([[10.5, 20.25, 30, 40], [0, 0, 0, 0], [0, 0, 0, 0]])
    r = driver.new_result(3, gt)
    r[1] = [11.123456789012345, 21.5, 30.25, 41.0]; r[2] = [1e-3, 2e5, 3.0, 4.0]
    assert r[0].tolist() == gt[0].tolist()
    driver.save_result(str(tmp_path / "a" / "t.txt"), r, "rgbt")
    driver.save_result(str(tmp_path / "a" / "e.txt"), r, "rgbe")
    np.savetxt(str(tmp_path / "ref_t.txt"), r)                                     # test_rgbt_mgpus.py:116
    np.savetxt(str(tmp_path / "ref_e.txt"), r, fmt="%.14f", delimiter=",")         # test_rgbe_mgpus.py:83
    assert open(tmp_path / "a" / "t.txt", "rb").read() == open(tmp_path / "ref_t.txt", "rb").read()
    assert open(tmp_path / "a" / "e.txt", "rb").read() == open(tmp_path / "ref_e.txt", "rb").read()
    assert driver.new_result(5, gt).shape == (5, 4)                               # length mismatch branch (:94-96)


def test_plan_groups():
    groups = driver.plan_groups([5, 50, 7, 49, 6], 2)
    assert groups == [[1, 3], [2, 4], [0]]
    assert driver.plan_groups([], 4) == []
