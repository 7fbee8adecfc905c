"""Parity protocol of the tracking forward -- TEST INFRASTRUCTURE ONLY.

`compare_frame` checks one teacher-forced frame of n sequences (both sides started from the same box and therefore
from the same u8 crop) against the oracle and classifies every (frame, sequence):

  CE kept sets   bit-exact, except where the fp32 reference's own scores of the disagreeing tokens lie within TAU_CE
                 (relative) of its keep threshold -- a *documented near-tie* (north_star);
  argmax         bit-exact, except where the reference's two largest Hann-windowed responses are within TAU_AM;
  score map      |gpu - ref| <= 1e-2 of the map maximum;
  box            |gpu - ref| <= 0.5 px (image pixels, after map_box_back / clip_box).

A near-tie excuses ONLY the discrete decision it concerns.  When the CE sets differ at a near-tie the remaining stages
are still verified: the oracle is re-run with the implementation's kept sets forced (`forced_removed`) and score map,
argmax and box must agree with THAT output.  So every frame ends in one of
  'ok'          everything agrees with the reference directly
  'ok_ce_tie'   CE sets differ only at documented near-ties and everything downstream agrees with the reference
                arithmetic evaluated on the implementation's kept set
  'ok_am_tie'   score map within tolerance, argmax differs at a documented near-tie (box therefore not compared)
  'fail'        anything else
and the north-star statistic `within_0.5px` is also reported over ALL frames with no exemption at all.
"""
import numpy as np
import torch

from . import vipt_oracle as vo

# Relative distance of a reference CE score from the keep threshold below which keeping / dropping that token is a
# documented near-tie.  Tied to the error level of 16-bit GEMM operands on each weight set: the reference's own
# arithmetic with fp16-rounded operands (vipt_oracle.OPERAND_ROUND) moves the scores of the boundary tokens by up to
# 2.3e-4 (flat) / 6.3e-3 (peaked) on 64 frames; the engine measured over 2 100 frames (profiles/r2_parity_report.json,
# `ce_boundary_score_rel_err`): p99 3.0e-4..5.4e-4, max 6.4e-4 (flat); p99 4.2e-3..8.2e-3, max 1.3e-2 (peaked).
# TAU_CE ~ 1.5 x the largest error seen; tests/test_gpu_parity.py asserts max error < TAU_CE <= 6 x the measured p99,
# so the exemption cannot be wider than the demonstrated arithmetic error.
TAU_CE_BY_SET = {"flat": 1e-3, "peaked": 2e-2, "outlier": 2e-2}
TAU_AM = 2e-3     # relative gap of the reference's top-2 windowed responses
TOL_MAP = 1e-2
TOL_BOX = 0.5
N_S = (256, 180, 126)        # search tokens entering the three CE stages (attn_blocks.py:40, ceil(0.7 n))
KEEP = (180, 126, 89)


def gpu_stage_index(ce_score, removed):
    """raster position of every search slot entering each CE stage, reconstructed from the implementation's own
    scores (descending, ties -> lower slot first: ce_topk_kernel) and cross-checked against its `removed` lists.
    ce_score (3, n, 256) float, removed (3, n, 256) int -> list of three (n, N_S[s]) int64 arrays."""
    n = ce_score.shape[1]
    gi = np.tile(np.arange(256, dtype=np.int64), (n, 1))
    out = []
    for s in range(3):
        out.append(gi)
        order = np.argsort(-ce_score[s, :, :N_S[s]], axis=1, kind="stable")
        nxt = np.take_along_axis(gi, order[:, :KEEP[s]], 1)
        rem = np.take_along_axis(gi, order[:, KEEP[s]:], 1)
        if not np.array_equal(rem, removed[s, :, :N_S[s] - KEEP[s]]):
            raise AssertionError("stage %d: `removed` is not the score-descending tail of the implementation's own "
                                 "scores (ce_topk_kernel ordering)" % s)
        gi = nxt
    return out


def _ce_compare(ora, gpu_removed, gpu_score, gpu_index, b, tau_ce):
    """-> (status, first differing stage or -1, per-stage boundary gaps, per-stage score errors or nan)."""
    gaps, errs = [], []
    status, first = "equal", -1
    for s in range(3):
        sc = ora["ce_scores"][s][b].numpy().astype(np.float64)
        srt = np.sort(sc)[::-1]
        thr = 0.5 * (srt[KEEP[s] - 1] + srt[KEEP[s]])
        gaps.append(float((srt[KEEP[s] - 1] - srt[KEEP[s]]) / max(srt[KEEP[s] - 1], 1e-300)))
        if status != "equal":
            errs.append(float("nan"))
            continue
        # implementation scores on the same tokens (both sides hold the same candidates while the sets agree)
        oi = ora["ce_index"][s][b].numpy()
        pos_g = np.empty(256, np.float64); pos_g[:] = np.nan
        pos_g[gpu_index[s][b]] = gpu_score[s, b, :N_S[s]]
        g = pos_g[oi]
        near = np.argsort(np.abs(sc - thr))[:16]                 # the 16 tokens closest to the keep threshold
        errs.append(float(np.max(np.abs(g[near] - sc[near]) / np.maximum(sc[near], 1e-300))))
        a = set(int(v) for v in gpu_removed[s, b, :N_S[s] - KEEP[s]])
        r = set(int(v) for v in ora["removed_indexes_s"][s][b].tolist())
        if a != r:
            first = s
            tok = np.array(sorted(a ^ r))
            col = {int(p): k for k, p in enumerate(oi)}
            d = np.array([abs(sc[col[int(p)]] - thr) / thr for p in tok])
            status = "near_tie" if d.max() < tau_ce else "mismatch"
    return status, first, gaps, errs


def _downstream(score_map_g, box_g, best_g, ref_out, k, ref_box, window):
    """score map / argmax / box / best score of one sequence against row k of a reference forward output."""
    sm_r = ref_out["score_map"][k].reshape(-1).numpy()
    sm_err = float(np.abs(score_map_g - sm_r).max() / sm_r.max())
    resp_r = ref_out["response"][k].reshape(-1).numpy()
    top = np.sort(resp_r)[::-1]
    am_gap = float((top[0] - top[1]) / top[0])
    am_g = int(np.argmax(score_map_g * window))      # argmax of the implementation's own windowed response
    am_eq = am_g == int(np.argmax(resp_r))
    box_err = float(np.abs(np.asarray(box_g) - np.asarray(ref_box)).max())
    best_err = float(abs(best_g - float(ref_out["best"][k])) / float(ref_out["best"][k]))
    return sm_err, am_eq, am_gap, box_err, best_err


def compare_frame(oracle, gpu, n=None, tau_ce=TAU_CE_BY_SET["flat"]):
    """oracle: OracleBatchTracker right after track(); gpu: dict(boxes (B,4), scores (B,), score_map (B,256),
    removed (3,B,256), ce_score (3,B,256)) of the same frame, teacher-forced.  -> list of n per-sequence records."""
    last = oracle.last
    ora = last["out"]
    n = oracle.n if n is None else n
    removed = np.asarray(gpu["removed"])[:, :n]
    ce_score = np.asarray(gpu["ce_score"])[:, :n]
    gidx = gpu_stage_index(ce_score, removed)
    window = oracle.window.reshape(-1).numpy()
    recs = []
    redo = []
    for b in range(n):
        st, first, gaps, errs = _ce_compare(ora, removed, ce_score, gidx, b, tau_ce)
        rec = {"ce_status": st, "ce_first_diff": first, "ce_gap": gaps, "ce_err": errs}
        sm_err, am_eq, am_gap, box_err, best_err = _downstream(
            np.asarray(gpu["score_map"])[b].reshape(-1), gpu["boxes"][b], float(gpu["scores"][b]), ora, b,
            oracle.state[b], window)
        rec.update(sm_err=sm_err, argmax_equal=am_eq, argmax_gap=am_gap, box_err=box_err, best_err=best_err)
        rec["ref_tie"] = bool(min(gaps) < tau_ce or am_gap < TAU_AM)
        if st == "near_tie":
            redo.append(b)
        recs.append(rec)
    if redo:
        # reference arithmetic on the implementation's kept sets (forced removal), for the sequences at a CE near-tie
        forced = [torch.from_numpy(removed[s, redo, :N_S[s] - KEEP[s]].astype(np.int64)) for s in range(3)]
        saved = oracle.state
        oracle.state = last["prev_state"]
        fo = oracle.forward(last["x"], forced_removed=forced, rows=redo)
        fboxes = oracle.boxes_from(fo, last["rfs"], rows=redo)
        oracle.state = saved
        for k, b in enumerate(redo):
            sm_err, am_eq, am_gap, box_err, best_err = _downstream(
                np.asarray(gpu["score_map"])[b].reshape(-1), gpu["boxes"][b], float(gpu["scores"][b]), fo, k, fboxes[k],
                window)
            recs[b]["forced"] = dict(sm_err=sm_err, argmax_equal=am_eq, argmax_gap=am_gap, box_err=box_err,
                                     best_err=best_err)
    for rec in recs:
        rec["verdict"] = _verdict(rec)
    return recs


def _downstream_ok(d):
    if d["sm_err"] > TOL_MAP or d["best_err"] > TOL_MAP:
        return "fail"
    if d["argmax_equal"]:
        return "ok" if d["box_err"] <= TOL_BOX else "fail"
    return "ok_am_tie" if d["argmax_gap"] < TAU_AM else "fail"


def _verdict(rec):
    if rec["ce_status"] == "mismatch":
        return "fail"
    if rec["ce_status"] == "equal":
        return _downstream_ok(rec)
    v = _downstream_ok(rec["forced"])
    return "fail" if v == "fail" else ("ok_ce_tie" if v == "ok" else v)


def summarise(recs):
    """recs: flat list of per-(frame, sequence) records -> report dict (fractions, quantiles)."""
    n = len(recs)
    q = lambda a, p: float(np.quantile(np.asarray(a, np.float64), p)) if len(a) else None
    verd = {}
    for r in recs:
        verd[r["verdict"]] = verd.get(r["verdict"], 0) + 1
    non_tie = [r for r in recs if not r["ref_tie"]]
    ce_equal = [r for r in recs if r["ce_status"] == "equal"]
    ce_err = [[r["ce_err"][s] for r in recs if not np.isnan(r["ce_err"][s])] for s in range(3)]
    return {
        "frames": n,
        "verdicts": verd,
        "fail": verd.get("fail", 0),
        "ref_near_tie_rate": 1.0 - len(non_tie) / n,
        "ce_status": {k: sum(r["ce_status"] == k for r in recs) for k in ("equal", "near_tie", "mismatch")},
        "ce_first_diff_stage": {str(s): sum(r["ce_first_diff"] == s for r in recs) for s in range(3)},
        "ce_boundary_score_rel_err": [{"p50": q(e, 0.5), "p99": q(e, 0.99), "max": q(e, 1.0)} for e in ce_err],
        "argmax_equal_all": sum(r["argmax_equal"] for r in recs) / n,
        "argmax_equal_ce_equal": (sum(r["argmax_equal"] for r in ce_equal) / len(ce_equal)) if ce_equal else None,
        "score_map_err_ce_equal": {"p50": q([r["sm_err"] for r in ce_equal], 0.5), "p99": q([r["sm_err"] for r in ce_equal], 0.99),
                                   "max": q([r["sm_err"] for r in ce_equal], 1.0)},
        "score_map_err_all": {"p50": q([r["sm_err"] for r in recs], 0.5), "p99": q([r["sm_err"] for r in recs], 0.99),
                              "max": q([r["sm_err"] for r in recs], 1.0)},
        "score_map_err_vs_forced": {"max": q([r["forced"]["sm_err"] for r in recs if "forced" in r], 1.0),
                                    "n": sum("forced" in r for r in recs)},
        "within_0.5px_all": sum(r["box_err"] <= TOL_BOX for r in recs) / n,
        "within_0.5px_non_tie": (sum(r["box_err"] <= TOL_BOX for r in non_tie) / len(non_tie)) if non_tie else None,
        "within_0.5px_ce_equal": (sum(r["box_err"] <= TOL_BOX for r in ce_equal) / len(ce_equal)) if ce_equal else None,
        "within_0.5px_verified": sum((r["forced"]["box_err"] if "forced" in r else r["box_err"]) <= TOL_BOX
                                     for r in recs) / n,
        "box_err_px": {"p50": q([r["box_err"] for r in recs], 0.5), "p99": q([r["box_err"] for r in recs], 0.99),
                       "max": q([r["box_err"] for r in recs], 1.0)},
    }
