"""Tracker-level behaviour of the ViPTTrack.initialize/track mirror: batched == single sequence, the reference's error
behaviour, the host-frame upload paths, the checkpoint path.  Parity against the oracle (teacher-forced and
free-running, all BASELINE.json configurations) lives in tests/test_gpu_parity.py."""
import os

import numpy as np
import pytest
import torch

from vipt_b200 import synth
from vipt_b200.tracker import BatchedViPTTracker, ViPTTrack

pytestmark = pytest.mark.gpu


class _Params:
    pass


def _params(sd, ptype):
    p = _Params()
    p.cfg = {"TRAIN": {"PROMPT": {"TYPE": ptype}}}
    p.state_dict = sd
    p.template_factor, p.template_size, p.search_factor, p.search_size = 2.0, 128, 4.0, 256
    p.save_all_boxes = False
    return p


def test_batched_matches_single():
    """8 sequences in one batch give the boxes of 8 single-sequence trackers (graph replay included)."""
    ptype = "vipt_deep"
    sd = synth.synthetic_state_dict(ptype, 0)
    seqs = [synth.synthetic_sequence(20 + i, 4, "rgbt") for i in range(8)]
    bt = BatchedViPTTracker(sd, 8, ptype, use_graph=True)
    bt.initialize([s[0][0] for s in seqs], [s[1] for s in seqs])
    batched = [bt.track([s[0][t] for s in seqs])[0].copy() for t in range(1, 4)]
    st = BatchedViPTTracker(sd, 1, ptype, use_graph=False)
    for i in (0, 5):
        st.initialize([seqs[i][0][0]], [seqs[i][1]])
        for t in range(1, 4):
            b, _ = st.track([seqs[i][0][t]])
            assert np.abs(b[0] - batched[t - 1][i]).max() < 1e-6, (i, t)


def test_too_small_box_raises():
    sd = synth.synthetic_state_dict("vipt_shaw", 0)
    trk = BatchedViPTTracker(sd, 1, "vipt_shaw")
    frames, _, _ = synth.synthetic_sequence(1, 1, "rgbe")
    with pytest.raises(Exception, match="Too small bounding box"):
        trk.initialize([frames[0]], [[5.0, 5.0, 0.0, 3.0]])


@pytest.mark.parametrize("pinned", [False, True])
def test_host_roi_upload_matches_device_frames(pinned):
    """track(host frames) uploads only the search rectangle (csrc/engine3.cu); the boxes must be bit-identical
    to the path that keeps whole frames on the device, including crops that overhang the frame.
    pinned=False: pageable ndarrays, one 2-D copy per sequence; pinned=True: page-locked tensors, one zero-copy gather
    kernel per step (the third frame is an odd-offset view, so source and destination alignments differ)."""
    B, T = 3, 6
    sd = synth.synthetic_state_dict("vipt_deep", 0)
    seqs = [synth.synthetic_sequence(40 + i, T, "rgbt") for i in range(B)]
    if pinned:
        def pin(f, odd):
            flat = torch.empty(f.size + 16, dtype=torch.uint8).pin_memory()
            v = flat[3:3 + f.size] if odd else flat[:f.size]
            v.copy_(torch.from_numpy(f).reshape(-1))
            return v.view(f.shape)
        seqs = [([pin(f, i == 2) for f in s[0]], s[1]) for i, s in enumerate(seqs)]
        as_dev = lambda f: f.cuda()
    else:
        as_dev = lambda f: torch.from_numpy(f).cuda()
    H, W = seqs[0][0][0].shape[:2]
    boxes0 = [list(seqs[0][1]), [2.0, 3.0, 50.0, 40.0], [W - 45.0, H - 38.0, 60.0, 50.0]]   # interior, two corners
    a = BatchedViPTTracker(sd, B, "vipt_deep", use_graph=False)
    b = BatchedViPTTracker(sd, B, "vipt_deep", use_graph=False)
    a.initialize([s[0][0] for s in seqs], boxes0)
    b.initialize([as_dev(s[0][0]) for s in seqs], boxes0)
    for t in range(1, T):
        ba, sa = a.track([s[0][t] for s in seqs])
        bb, sb = b.track([as_dev(s[0][t]) for s in seqs])
        assert 0 < a.h2d_bytes < B * H * W * 6
        np.testing.assert_array_equal(ba, bb)
        np.testing.assert_array_equal(sa, sb)
    # two steps enqueued without reading results in between
    a.track_async([s[0][1] for s in seqs]); a.track_async([s[0][2] for s in seqs])
    b.track_async([as_dev(s[0][1]) for s in seqs])
    b1 = b.engine.results()[0]
    b.track_async([as_dev(s[0][2]) for s in seqs])
    b2 = b.engine.results()[0]
    if pinned:       # rectangles evaluated on the device from the current state: both steps stay partial uploads
        a1, _ = a.collect()
        assert 0 < a.h2d_bytes < B * H * W * 6
        a2, _ = a.collect()
        assert 0 < a.h2d_bytes < B * H * W * 6
        np.testing.assert_array_equal(a1, b1)
        np.testing.assert_array_equal(a2, b2)
        with pytest.raises(RuntimeError):
            a.engine.collect()                   # nothing left to collect
    else:            # stale host mirror: the second step falls back to whole frames
        assert a.h2d_bytes == B * H * W * 6
        np.testing.assert_array_equal(a.engine.results()[0], b2)


def test_checkpoint_file_path(tmp_path):
    """ViPTTrack(params) with params.checkpoint -> torch.load(...)['net'] (ViPT/lib/test/tracker/vipt.py:21), the
    way the reference is always constructed; same boxes as the in-memory state_dict path."""
    sd = synth.weight_set("peaked", "vipt_deep")
    path = str(tmp_path / "ViPT_deep_rgbd.pth.tar")
    torch.save({"net": sd, "epoch": 60}, path)
    frames, box0, _ = synth.synthetic_sequence(31, 4, "rgbd")
    p = _params(None, "vipt_deep")
    p.state_dict = None
    p.checkpoint = path
    a = ViPTTrack(p)
    b = ViPTTrack(_params(sd, "vipt_deep"))
    a.initialize(frames[0], {"init_bbox": list(box0)})
    b.initialize(frames[0], {"init_bbox": list(box0)})
    for t in range(1, 4):
        ra, rb = a.track(frames[t]), b.track(frames[t])
        assert ra["target_bbox"] == rb["target_bbox"] and ra["best_score"] == rb["best_score"]


def test_frames_produced_on_another_stream_are_ordered():
    """Device frames are produced on the caller's stream while the engine runs on its own: a slow producer
    (a long kernel queued before the frame is written) must not let the crop kernel read the frame too early."""
    sd = synth.weight_set("peaked", "vipt_deep")
    seq, box0, _ = synth.synthetic_sequence(33, 3, "rgbt")
    dev = torch.device("cuda")
    ref = BatchedViPTTracker(sd, 1, "vipt_deep", use_graph=False)
    ref.initialize([torch.from_numpy(seq[0]).to(dev)], [box0])
    want = [ref.track([torch.from_numpy(seq[t]).to(dev)])[0].copy() for t in (1, 2)]
    trk = BatchedViPTTracker(sd, 1, "vipt_deep", use_graph=False)
    trk.initialize([torch.from_numpy(seq[0]).to(dev)], [box0])
    big = torch.randn(8192, 8192, device=dev)
    host = [torch.from_numpy(seq[t]).pin_memory() for t in (1, 2)]
    for k in range(2):
        frame = torch.zeros(seq[0].shape, dtype=torch.uint8, device=dev)      # stale contents: all black
        for _ in range(6):
            big = big @ big * 1e-4          # ~ 40 ms of work on the current stream ahead of the frame's producer
        frame.copy_(host[k], non_blocking=True)
        got, _ = trk.track([frame])
        np.testing.assert_array_equal(got, want[k])


@pytest.mark.parametrize("switch", ["VIPT_NO_ZEROCOPY", "VIPT_NO_SPECULATION"])
def test_host_path_switches(monkeypatch, switch):
    """The two host-frame switches (read once when the engine is created): page-locked frames uploaded by per-sequence
    2-D copies instead of the zero-copy gather kernel / no speculative copy-engine upload next to the running step.
    Same boxes as the device-frame path, pipelined call pattern included."""
    monkeypatch.setenv(switch, "1")
    B, T = 2, 5
    sd = synth.weight_set("peaked", "vipt_deep")
    seqs = [synth.synthetic_sequence(60 + i, T, "rgbt") for i in range(B)]
    pinned = [[torch.from_numpy(f).pin_memory() for f in s[0]] for s in seqs]
    a = BatchedViPTTracker(sd, B, "vipt_deep", use_graph=False)
    monkeypatch.delenv(switch)
    b = BatchedViPTTracker(sd, B, "vipt_deep", use_graph=False)
    a.initialize([p[0] for p in pinned], [s[1] for s in seqs])
    b.initialize([torch.from_numpy(s[0][0]).cuda() for s in seqs], [s[1] for s in seqs])
    want = [b.track([torch.from_numpy(s[0][t]).cuda() for s in seqs])[0].copy() for t in range(1, T)]
    a.track_async([p[1] for p in pinned])
    got = []
    for t in range(2, T):                      # one step in flight: enqueue t, collect t - 1
        a.track_async([p[t] for p in pinned])
        got.append(a.collect()[0].copy())
    got.append(a.collect()[0].copy())
    for g, w in zip(got, want):
        np.testing.assert_array_equal(g, w)
