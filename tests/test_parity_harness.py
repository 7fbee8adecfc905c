"""CPU tests of the parity infrastructure (no GPU):
  * the oracle on the "peaked" weight set is pinned against the UNMODIFIED reference (tests/golden/peaked_golden.npz:
    model forward of deep + shaw, and the 50-frame deep-RGBD run of BASELINE.json configs[0]);
  * the batched oracle tracker equals the single-sequence one;
  * forced removal with the reference's own decision reproduces the reference output bit for bit;
  * the comparator (oracle/parity.py) accepts a stand-in implementation with fp16-rounded operands
    (vipt_oracle.OPERAND_ROUND) and rejects a broken one (a wrong CE decision far from any tie, a shifted box).
"""
import os

import numpy as np
import pytest
import torch

from oracle import batch, parity, vipt_oracle as vo
from vipt_b200 import synth


@pytest.mark.parametrize("name,ptype", [("deep", "vipt_deep"), ("shaw", "vipt_shaw")])
def test_peaked_forward_matches_reference_golden(golden_dir, name, ptype):
    g = np.load(os.path.join(golden_dir, "peaked_golden.npz"))
    sd = synth.weight_set("peaked", ptype)
    gen = torch.Generator().manual_seed(321)
    z = torch.randn(4, 6, 128, 128, generator=gen); x = torch.randn(4, 6, 256, 256, generator=gen)
    with torch.no_grad():
        o = vo.model_forward(sd, z, x, vo.ce_template_mask(4), ptype)
    for k in ("pred_boxes", "score_map", "size_map", "offset_map"):
        np.testing.assert_allclose(o[k].numpy(), g[name + "_" + k], rtol=0, atol=5e-5)
    for i in range(3):
        assert np.array_equal(o["removed_indexes_s"][i].numpy(), g[name + "_removed_%d" % i])
    np.testing.assert_allclose(o["attn"][:, :, 27, :].numpy(), g[name + "_attn_row27"], rtol=1e-4, atol=2e-6)


def test_peaked_50_frame_run_matches_reference_golden(golden_dir):
    """BASELINE.json configs[0]: the oracle's free run over 50 frames against the unmodified reference's boxes."""
    g = np.load(os.path.join(golden_dir, "peaked_golden.npz"))
    T = int(g["track_T"])
    assert T == 51
    frames, box0, _ = synth.synthetic_sequence(int(g["track_seed"]), T, "rgbd")
    tr = batch.OracleBatchTracker(synth.weight_set("peaked", "vipt_deep"), "vipt_deep", use_cv2=False)
    tr.initialize([frames[0]], [box0])
    worst = 0.0
    for t in range(1, T):
        b, s = tr.track([frames[t]])
        worst = max(worst, float(np.abs(b[0] - g["track_boxes"][t]).max()))
        assert abs(float(s[0]) - g["track_scores"][t]) < 2e-4, t
    assert worst < 2e-2, worst          # fp32 summation-order noise only; no drift over 50 frames


def test_batched_oracle_equals_single():
    sd = synth.weight_set("peaked", "vipt_deep")
    seqs = [synth.synthetic_sequence(300 + i, 3, "rgbt") for i in range(3)]
    bt = batch.OracleBatchTracker(sd, "vipt_deep")
    bt.initialize([s[0][0] for s in seqs], [s[1] for s in seqs])
    st = vo.OracleTracker(sd, "vipt_deep")
    st.initialize(seqs[1][0][0], {"init_bbox": list(seqs[1][1])})
    for t in (1, 2):
        b, s = bt.track([q[0][t] for q in seqs])
        r = st.track(seqs[1][0][t])
        np.testing.assert_allclose(b[1], r["target_bbox"], atol=2e-3)
        assert abs(float(s[1]) - r["best_score"]) < 1e-5


def _as_gpu(out, boxes, scores, n):
    rem = np.zeros((3, n, 256), np.int64); sc = np.zeros((3, n, 256), np.float32)
    for s in range(3):
        rem[s, :, :out["removed_indexes_s"][s].shape[1]] = out["removed_indexes_s"][s].numpy()
        sc[s, :, :out["ce_scores"][s].shape[1]] = out["ce_scores"][s].numpy()
    return {"boxes": boxes, "scores": scores, "score_map": out["score_map"].reshape(n, 256).numpy(), "removed": rem,
            "ce_score": sc}


def test_forced_removal_and_comparator():
    n = 4
    sd = synth.weight_set("peaked", "vipt_deep")
    tau = parity.TAU_CE_BY_SET["peaked"]
    seqs = [synth.synthetic_sequence(300 + i, 2, "rgbt") for i in range(n)]
    ora = batch.OracleBatchTracker(sd, "vipt_deep"); ora.initialize([s[0][0] for s in seqs], [s[1] for s in seqs])
    emu = batch.OracleBatchTracker(sd, "vipt_deep"); emu.initialize([s[0][0] for s in seqs], [s[1] for s in seqs])
    fr = [s[0][1] for s in seqs]
    ob, osc = ora.track(fr)
    out = ora.last["out"]
    # (1) forcing the reference's own decision changes nothing
    saved = ora.state; ora.state = ora.last["prev_state"]
    fo = ora.forward(ora.last["x"], forced_removed=out["removed_indexes_s"])
    ora.state = saved
    assert torch.equal(fo["score_map"], out["score_map"])
    for s in range(3):
        assert torch.equal(fo["removed_indexes_s"][s], out["removed_indexes_s"][s])
    # (2) the reference against itself: all 'ok', zero error
    recs = parity.compare_frame(ora, _as_gpu(out, ob, osc, n), tau_ce=tau)
    assert [r["verdict"] for r in recs] == ["ok"] * n and max(r["box_err"] for r in recs) == 0.0
    # (3) a stand-in implementation with fp16-rounded GEMM / attention operands passes the protocol
    vo.OPERAND_ROUND = lambda t: t.half().float()
    try:
        eb, esc = emu.track(fr)
    finally:
        vo.OPERAND_ROUND = None
    recs = parity.compare_frame(ora, _as_gpu(emu.last["out"], eb, esc, n), tau_ce=tau)
    assert all(r["verdict"] in ("ok", "ok_ce_tie", "ok_am_tie") for r in recs), [r["verdict"] for r in recs]
    assert max(e for r in recs for e in r["ce_err"] if not np.isnan(e)) < tau
    # (4) broken implementations are rejected: the best-ranked token dropped instead of the boundary one ...
    forced0 = out["removed_indexes_s"][0].clone()
    forced0[0, 0] = int(out["ce_index"][0][0][int(np.argmax(out["ce_scores"][0][0].numpy()))])
    saved = ora.state; ora.state = ora.last["prev_state"]
    wrong = ora.forward(ora.last["x"], forced_removed=[forced0, None, None])
    wb = ora.boxes_from(wrong, ora.last["rfs"])
    ora.state = saved
    bad = _as_gpu(wrong, wb, wrong["best"].numpy(), n)
    sc0 = out["ce_scores"][0][0].numpy()
    low, nxt = int(out["removed_indexes_s"][0][0, 0]), int(out["removed_indexes_s"][0][0, 1])
    bad["ce_score"][0, 0, int(forced0[0, 0])] = 0.5 * (sc0[low] + sc0[nxt])     # a score consistent with its own decision
    recs = parity.compare_frame(ora, bad, tau_ce=tau)
    assert recs[0]["verdict"] == "fail" and recs[0]["ce_status"] == "mismatch" and recs[1]["verdict"] == "ok"
    # ... and a box off by one pixel with everything else right
    bad = _as_gpu(out, ob.copy(), osc.copy(), n)
    bad["boxes"][2, 0] += 1.0
    recs = parity.compare_frame(ora, bad, tau_ce=tau)
    assert recs[2]["verdict"] == "fail" and recs[1]["verdict"] == "ok"
    s = parity.summarise(recs)
    assert s["fail"] == 1 and s["frames"] == n
