"""The parity protocol (oracle/parity.py) on BASELINE.json's configurations, on the GPU, through the
public tracker API.  Every frame of every sequence is classified; nothing is hand-picked and no share of frames may
fail.  The full-length run (>= 400 frames per configuration) is scripts/parity_report.py ->
profiles/r2_parity_report.json; these tests run the same code on fewer frames."""
import os

import numpy as np
import pytest

from oracle import parity, parity_run

pytestmark = pytest.mark.gpu


def _check(summary, recs):
    bad = [(r["frame"], r["seq"], r["ce_status"], r["sm_err"], r["box_err"]) for r in recs if r["verdict"] == "fail"]
    assert not bad, bad[:10]
    assert summary["ce_status"]["mismatch"] == 0
    assert summary["replicas_bit_identical"]
    # frames whose CE sets equal the reference's: score map <= 1e-2, and the box within 0.5 px unless the argmax sits
    # at a documented near-tie -- at least 99.9 % (north_star) once those are set aside
    eq = [r for r in recs if r["ce_status"] == "equal" and r["verdict"] != "ok_am_tie"]
    assert eq and np.mean([r["box_err"] <= parity.TOL_BOX for r in eq]) >= 0.999
    assert max(r["sm_err"] for r in eq) <= parity.TOL_MAP
    # every frame, verified directly or against the reference arithmetic on the implementation's kept set
    assert summary["within_0.5px_verified"] >= 0.999 or summary["verdicts"].get("ok_am_tie", 0) > 0
    # the tie band is not wider than the demonstrated arithmetic error allows, and the error stays inside it
    errs = [e for r in recs for e in r["ce_err"] if not np.isnan(e)]
    assert max(errs) < summary["tau_ce"], (max(errs), summary["tau_ce"])
    assert summary["tau_ce"] <= 6.0 * np.quantile(errs, 0.99), (summary["tau_ce"], np.quantile(errs, 0.99))


def test_config0_50_frames_bs1(golden_dir):
    """BASELINE.json configs[0]: ViPT-deep RGBD, one 50-frame sequence, batch 1, CUDA-graph replay; teacher-forced
    against the oracle and free-running against the oracle and the UNMODIFIED reference's boxes."""
    g = np.load(os.path.join(golden_dir, "peaked_golden.npz"))
    ptype, modality, B = parity_run.CONFIGS["config0_deep_rgbd_bs1"]
    summary, recs = parity_run.run_config("peaked", ptype, modality, B, 1, 50, golden_boxes=g["track_boxes"])
    assert summary["frames"] == 50
    _check(summary, recs)
    fr = summary["free_running"]
    assert fr["oracle_vs_unmodified_reference_px_max"] < 2e-2
    assert fr["vs_unmodified_reference_px"]["frame1"] < 0.5
    print("free-running vs unmodified reference:", fr["vs_unmodified_reference_px"])


@pytest.mark.parametrize("name,weights,T", [
    ("config1_deep_rgbt_64", "peaked", 6),
    ("config1_deep_rgbt_64", "flat", 3),
    ("config2_deep_rgbe_64", "peaked", 3),
    ("config3_shaw_rgbd_32", "peaked", 3),
    ("config4_deep_rgbt_bs1", "peaked", 12),
    ("config4_deep_rgbt_128", "peaked", 3),
])
def test_config_batches(name, weights, T):
    """16 distinct sequences (replicated to the configuration's per-GPU batch) x T frames, every frame checked."""
    ptype, modality, B = parity_run.CONFIGS[name]
    summary, recs = parity_run.run_config(weights, ptype, modality, B, 16, T)
    assert summary["frames"] == min(16, B) * T
    _check(summary, recs)
