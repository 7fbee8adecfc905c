"""Tracker-level parity (ViPTTrack.initialize/track mirror) against the CPU oracle:
 (i) teacher-forced: both sides start every frame from the oracle's previous box;
 (ii) free-running drift over the same frames."""
import os

import numpy as np
import pytest
import torch

from oracle import near_tie, vipt_oracle as vo
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


@pytest.mark.parametrize("ptype,modality,seed", [("vipt_deep", "rgbd", 11), ("vipt_shaw", "rgbe", 5)])
def test_teacher_forced_single_sequence(ptype, modality, seed):
    T = 8
    sd = synth.synthetic_state_dict(ptype, 0)
    frames, box0, _ = synth.synthetic_sequence(seed, T, modality)
    ref = vo.OracleTracker(sd, ptype)
    ref.initialize(frames[0], {"init_bbox": list(box0)})
    trk = ViPTTrack(_params(sd, ptype))
    assert trk.initialize(frames[0], {"init_bbox": list(box0)}) is None
    good = 0
    for t in range(1, T):
        r = ref.track(frames[t])
        o = trk.track(frames[t])
        assert set(o) == {"target_bbox", "best_score"} and len(o["target_bbox"]) == 4
        d = np.abs(np.asarray(o["target_bbox"]) - np.asarray(r["target_bbox"])).max()
        sm = trk._impl.engine.buffer("score_map").cpu()
        sm_err = ((sm - ref.last["out"]["score_map"]).abs().max() / ref.last["out"]["score_map"].max()).item()
        ok = d < 0.5 and sm_err < 1e-2 and abs(o["best_score"] - r["best_score"]) < 1e-2 * r["best_score"]
        good += ok
        # a frame may disagree only where the fp32 reference is at a documented near-tie
        assert ok or bool(near_tie.is_near_tie(ref.last["out"], ref.window)[0]), \
            "frame %d: box %.3f px, score map %.2e, and the reference is not at a near-tie" % (t, d, sm_err)
        # teacher forcing: overwrite the device state with the oracle's box
        trk._impl.set_state([r["target_bbox"]])
    assert good >= (T - 1) // 2, "only %d of %d frames agree" % (good, T - 1)


def test_golden_track_free_running(golden_dir):
    """free-running against boxes produced by the unmodified reference (committed golden)."""
    g = np.load(os.path.join(golden_dir, "track_golden.npz"))
    name, ptype, modality = "deep_rgbd", "vipt_deep", "rgbd"
    T = int(g[name + "_T"])
    frames, box0, _ = synth.synthetic_sequence(int(g[name + "_seed"]), T, modality)
    trk = BatchedViPTTracker(synth.synthetic_state_dict(ptype, 0), 1, ptype)
    trk.initialize([frames[0]], [box0])
    drift = []
    for t in range(1, T):
        boxes, _ = trk.track([frames[t]])
        drift.append(float(np.abs(boxes[0] - g[name + "_boxes"][t]).max()))
    print("free-running drift per frame (px):", np.round(drift, 3).tolist())
    assert drift[0] < 0.5


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


@pytest.mark.parametrize("ptype,modality,B,checked", [
    ("vipt_deep", "rgbt", 64, (0, 1, 4)),       # BASELINE.json configs[1]: 64 LasHeR-shaped sequences per B200
    ("vipt_deep", "rgbe", 64, (2,)),            # configs[2]: VisEvent-shaped, the per-GPU shard of the 2/4/8-GPU runs
    ("vipt_shaw", "rgbd", 32, (12, 14, 15)),    # configs[3]: shallow prompt, 256 sequences over 8 GPUs = 32 per GPU
    ("vipt_deep", "rgbt", 128, (0, 1, 4)),      # configs[4]: 1024 sequences over 8 GPUs = 128 per GPU
])
def test_full_size_batch_against_oracle(ptype, modality, B, checked):
    """BASELINE.json configs at full per-GPU size: B concurrent sequences in one engine (the 2-CTA GEMM, persistent
    attention and cooperative prompt kernels are selected only at this scale), CUDA-graph replay.
    (i) slots carrying the same sequence give bit-identical boxes and scores (no cross-slot leakage, deterministic);
    (ii) slots are checked against the CPU oracle on the first tracked frame (both sides start from the init box):
         boxes within 0.5 px and score maps within 1e-2 of the map maximum unless the fp32 reference is at a near-tie.
    `checked` = the sequences of this seed range at which the fp32 reference is NOT at a documented near-tie on
    frame 1 (synthetic weights give flat CE scores: most sequences are)."""
    T, ND = 3, 16
    sd = synth.synthetic_state_dict(ptype, 0)
    base = [synth.synthetic_sequence(300 + i, T, modality) for i in range(ND)]
    trk = BatchedViPTTracker(sd, B, ptype, use_graph=True)
    trk.initialize([base[i % ND][0][0] for i in range(B)], [base[i % ND][1] for i in range(B)])
    boxes1, scores1 = trk.track([base[i % ND][0][1] for i in range(B)])
    boxes1, scores1 = boxes1.copy(), scores1.copy()
    sm1 = trk.engine.buffer("score_map").cpu().view(B, -1).clone()
    agree = 0
    for i in checked:
        ref = vo.OracleTracker(sd, ptype)
        ref.initialize(base[i][0][0], {"init_bbox": list(base[i][1])})
        r = ref.track(base[i][0][1])
        ref_sm = ref.last["out"]["score_map"].view(-1)
        tie = bool(near_tie.is_near_tie(ref.last["out"], ref.window)[0])
        for slot in (i, i + B - ND):
            d = float(np.abs(boxes1[slot] - np.asarray(r["target_bbox"])).max())
            e = float((sm1[slot] - ref_sm).abs().max() / ref_sm.max())
            ok = d < 0.5 and e < 1e-2 and abs(scores1[slot] - r["best_score"]) < 1e-2 * r["best_score"]
            agree += ok
            assert ok or tie, "slot %d: box %.3f px, score map %.2e, reference not at a near-tie" % (slot, d, e)
    assert 3 * agree >= 2 * 2 * len(checked), agree
    for t in range(1, T):
        if t > 1:
            boxes, scores = trk.track([base[i % ND][0][t] for i in range(B)])
        else:
            boxes, scores = boxes1, scores1
        for i in range(ND):
            for rep in range(1, B // ND):
                np.testing.assert_array_equal(boxes[i], boxes[i + rep * ND])
                assert scores[i] == scores[i + rep * ND]
        assert np.isfinite(boxes).all()
