"""Parity report of the tracking forward on BASELINE.json's five configurations (GPU box).

    python scripts/parity_report.py [--frames 25] [--out profiles/r2_parity_report.json] [--quick]

For every configuration: teacher-forced comparison of every frame against the CPU oracle (oracle/parity.py: CE kept
sets, argmax, score map <= 1e-2, box <= 0.5 px; near-ties excuse only the discrete decision they concern and are then
verified against the reference arithmetic on the implementation's kept set) and the free-running drift.
configs[0] additionally against the 50-frame run of the UNMODIFIED reference (tests/golden/peaked_golden.npz).
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

from oracle import parity_run  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=25)
    ap.add_argument("--distinct", type=int, default=16)
    ap.add_argument("--out", default=os.path.join(ROOT, "profiles", "r2_parity_report.json"))
    ap.add_argument("--quick", action="store_true", help="4 frames per configuration (smoke run)")
    ap.add_argument("--weights", default="peaked,flat")
    ap.add_argument("--only", default="")
    args = ap.parse_args()
    torch.set_num_threads(os.cpu_count() or 1)
    T = 4 if args.quick else args.frames
    golden = np.load(os.path.join(ROOT, "tests", "golden", "peaked_golden.npz"))
    report = {"protocol": "oracle/parity.py", "host_threads": os.cpu_count(), "configs": {}}
    plan = []
    for w in args.weights.split(","):
        for name, (ptype, modality, B) in parity_run.CONFIGS.items():
            if args.only and args.only not in name:
                continue
            if w == "flat" and name != "config1_deep_rgbt_64":
                continue                 # the flat (default-init) set only on the headline configuration
            frames = (10 if args.quick else 50) if B == 1 else T
            plan.append((w, name, ptype, modality, B, frames))
    for w, name, ptype, modality, B, frames in plan:
        t0 = time.time()
        gb = golden["track_boxes"] if (w == "peaked" and name == "config0_deep_rgbd_bs1") else None
        summary, recs = parity_run.run_config(w, ptype, modality, B, args.distinct, frames, golden_boxes=gb,
                                              log=lambda s: print(s, flush=True))
        summary["seconds"] = round(time.time() - t0, 1)
        summary["not_ok_frames"] = [{k: r[k] for k in ("frame", "seq", "verdict", "ce_status", "ce_first_diff", "sm_err",
                                                       "box_err", "argmax_equal", "argmax_gap")}
                                    for r in recs if r["verdict"] == "fail"][:20]
        report["configs"]["%s/%s" % (name, w)] = summary
        print(json.dumps({"%s/%s" % (name, w): {k: summary[k] for k in (
            "frames", "verdicts", "ce_status", "within_0.5px_all", "within_0.5px_ce_equal", "within_0.5px_verified",
            "ce_boundary_score_rel_err", "score_map_err_ce_equal", "seconds")}}), flush=True)
        with open(args.out, "w") as f:
            json.dump(report, f, indent=1)
    tot = sum(c["frames"] for c in report["configs"].values())
    fails = sum(c["fail"] for c in report["configs"].values())
    report["total_frames"] = tot
    report["total_fail"] = fails
    with open(args.out, "w") as f:
        json.dump(report, f, indent=1)
    print("parity report: %d frames, %d fail -> %s" % (tot, fails, args.out))
    return 0 if fails == 0 else 1


if __name__ == "__main__":
    sys.exit(main())
