"""Runs the parity protocol (oracle/parity.py) of one configuration on the GPU -- TEST INFRASTRUCTURE ONLY.

Used by tests/test_gpu_parity.py (asserts) and scripts/parity_report.py (writes profiles/r2_parity_report.json).
The engine under test is driven through the public tracker API (BatchedViPTTracker -> ctypes -> C ABI).

  (i)  teacher-forced: every frame both sides start from the oracle's previous box, so each frame is an independent
       comparison on identical crops;
  (ii) free-running: the engine runs from the initial box on its own; drift against the oracle's (and, where a golden
       exists, the unmodified reference's) trajectory.
"""
import time

import numpy as np
import torch

from . import parity
from .batch import OracleBatchTracker

# BASELINE.json configs: (prompt type, modality, sequences per GPU)
CONFIGS = {
    "config0_deep_rgbd_bs1": ("vipt_deep", "rgbd", 1),
    "config1_deep_rgbt_64": ("vipt_deep", "rgbt", 64),
    "config2_deep_rgbe_64": ("vipt_deep", "rgbe", 64),
    "config3_shaw_rgbd_32": ("vipt_shaw", "rgbd", 32),
    "config4_deep_rgbt_bs1": ("vipt_deep", "rgbt", 1),
    "config4_deep_rgbt_128": ("vipt_deep", "rgbt", 128),
}


def run_config(weights, ptype, modality, B, n_distinct, T, seed0=300, use_graph=True, free_run=True, golden_boxes=None,
               use_cv2=False, log=None):
    """-> (summary dict, flat list of per-(frame, sequence) records)."""
    from vipt_b200 import synth
    from vipt_b200.tracker import BatchedViPTTracker
    sd = synth.weight_set(weights, ptype)
    tau = parity.TAU_CE_BY_SET[weights]
    nd = min(n_distinct, B)
    seqs = [synth.synthetic_sequence(seed0 + i, T + 1, modality) for i in range(nd)]
    dev = torch.device("cuda", torch.cuda.current_device())
    devf = [[torch.from_numpy(s[0][t]).to(dev) for s in seqs] for t in range(T + 1)]
    rep = lambda per_seq: [per_seq[i % nd] for i in range(B)]
    box0 = [s[1] for s in seqs]

    ora = OracleBatchTracker(sd, ptype, use_cv2=use_cv2)
    ora.initialize([s[0][0] for s in seqs], box0)
    trk = BatchedViPTTracker(sd, B, ptype, use_graph=use_graph)
    eng = trk.engine
    trk.initialize(rep(devf[0]), rep(box0))
    recs, ora_boxes = [], [np.asarray(box0, np.float64)]
    t_or = 0.0
    replicas_identical = True
    for t in range(1, T + 1):
        t0 = time.time()
        ob, _ = ora.track([s[0][t] for s in seqs])
        t_or += time.time() - t0
        gb, gs = trk.track(rep(devf[t]))
        gpu = {"boxes": gb, "scores": gs, "score_map": eng.buffer("score_map").cpu().view(B, 256).numpy(),
               "removed": eng.buffer("removed").cpu().numpy(), "ce_score": eng.buffer("ce_score").cpu().numpy()}
        for r in range(1, B // nd):      # slots carrying the same sequence: bit-identical (no cross-slot leakage)
            replicas_identical &= bool(np.array_equal(gb[:nd], gb[r * nd:(r + 1) * nd]) and
                                       np.array_equal(gs[:nd], gs[r * nd:(r + 1) * nd]))
        fr = parity.compare_frame(ora, gpu, n=nd, tau_ce=tau)
        for i, r in enumerate(fr):
            r["frame"], r["seq"] = t, i
        recs += fr
        ora_boxes.append(ob.copy())
        trk.set_state(np.asarray(rep(list(ob))))      # teacher forcing
        if log:
            log("  frame %d/%d  verdicts %s" % (t, T, [r["verdict"] for r in fr][:8]))
    summary = parity.summarise(recs)
    summary.update(weights=weights, prompt_type=ptype, modality=modality, sequences_in_engine=B, distinct_sequences=nd,
                   frames_per_sequence=T, tau_ce=tau, tau_am=parity.TAU_AM, cuda_graph=bool(use_graph),
                   replicas_bit_identical=replicas_identical, oracle_seconds=round(t_or, 1))
    if free_run:
        trk.initialize(rep(devf[0]), rep(box0))
        drift = []
        gbs = []
        for t in range(1, T + 1):
            gb, _ = trk.track(rep(devf[t]))
            gbs.append(gb[:nd].copy())
            drift.append(np.abs(gb[:nd] - ora_boxes[t]).max(axis=1))
        drift = np.asarray(drift)                      # (T, nd)
        fr_sum = {"frames": [int(t) for t in (1, 5, 10, 25, 50) if t <= T]}
        fr_sum["drift_px_median"] = [float(np.median(drift[t - 1])) for t in fr_sum["frames"]]
        fr_sum["drift_px_max"] = [float(np.max(drift[t - 1])) for t in fr_sum["frames"]]
        fr_sum["sequences_within_0.5px_at_end"] = float(np.mean(drift[-1] <= 0.5))
        fr_sum["sequences_within_2px_at_end"] = float(np.mean(drift[-1] <= 2.0))
        fr_sum["frames_within_0.5px"] = float(np.mean(drift <= 0.5))
        if golden_boxes is not None:
            g = np.abs(np.asarray(gbs)[:, 0] - np.asarray(golden_boxes)[1:T + 1]).max(axis=1)
            fr_sum["vs_unmodified_reference_px"] = {"frame1": float(g[0]), "median": float(np.median(g)),
                                                    "max": float(g.max()), "at_end": float(g[-1]),
                                                    "frames_within_0.5px": float(np.mean(g <= 0.5))}
            o = np.abs(np.asarray(ora_boxes)[1:, 0] - np.asarray(golden_boxes)[1:T + 1]).max(axis=1)
            fr_sum["oracle_vs_unmodified_reference_px_max"] = float(o.max())
        summary["free_running"] = fr_sum
    return summary, recs
