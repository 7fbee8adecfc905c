"""Generate the committed golden vectors by running the UNMODIFIED reference here (build container).

    python tests/golden/make_golden.py

Needs /root/reference (see oracle/ref_harness.py); the GPU box never runs this.  Outputs:
  tests/golden/crop_golden.npz    cv2-backed sample_target + PreprocessorMM outputs
  tests/golden/model_golden.npz   build_viptrack(...).forward on seeded inputs (deep + shaw)
  tests/golden/track_golden.npz   ViPTTrack.initialize/track over seeded synthetic sequences
  tests/golden/frames_golden.npz  get_x_frame ('rgbrgb', 'rgbcolormap' with / without depth clip) on synthetic PNGs
  tests/golden/peaked_golden.npz  the "peaked" weight set (synth.WEIGHT_SETS): model forward (deep + shaw) and the
                                  50-frame deep-RGBD tracking run of BASELINE.json configs[0]  (--peaked-only)
Weights are not stored: vipt_b200.synth.synthetic_state_dict(seed) regenerates them.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ref_harness as rh  # noqa: E402
from vipt_b200 import synth  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def crop_golden():
    rh.import_reference()
    from lib.train.data.processing_utils import sample_target
    from lib.test.tracker.data_utils import PreprocessorMM
    pre = PreprocessorMM()
    frames, _, _ = synth.synthetic_sequence(7, 1, "rgbt")
    im = frames[0]
    H, W = im.shape[:2]
    cases = [  # box [x,y,w,h], factor, out size -- interior, overhang on every side, 2x, upscale, tiny
        ([250.0, 190.0, 60.0, 48.0], 4.0, 256), ([250.0, 190.0, 60.0, 48.0], 2.0, 128),
        ([-20.5, -10.25, 90.0, 70.0], 4.0, 256), ([W - 40.0, H - 30.0, 80.0, 64.0], 4.0, 256),
        ([300.0, 200.0, 128.0, 128.0], 4.0, 256),      # crop 512 -> exact 2x
        ([300.0, 200.0, 128.0, 128.0], 2.0, 128),      # crop 256 -> exact 2x (template)
        ([310.2, 222.7, 12.0, 10.0], 4.0, 256),        # upscale
        ([100.5, 50.5, 200.0, 300.0], 4.0, 256),       # crop larger than the image
        ([301.5, 201.5, 33.0, 31.0], 2.0, 128), ([10.0, 400.0, 63.9, 64.1], 4.0, 256),
    ]
    out = {"frame_seed": np.int64(7), "n": np.int64(len(cases))}
    for i, (box, f, s) in enumerate(cases):
        patch, rf, _ = sample_target(im, box, f, output_sz=s)
        t = pre.process(patch)
        out["box_%d" % i] = np.array(box); out["factor_%d" % i] = np.float64(f); out["size_%d" % i] = np.int64(s)
        out["patch_%d" % i] = patch; out["rf_%d" % i] = np.float64(rf)
        out["norm_sum_%d" % i] = t.double().sum().numpy(); out["norm_abs_%d" % i] = t.double().abs().sum().numpy()
    np.savez_compressed(os.path.join(OUT, "crop_golden.npz"), **out)


def model_golden():
    out = {}
    for name, yaml_name, ptype in (("deep", "deep_rgbt", "vipt_deep"), ("shaw", "shaw_rgbd", "vipt_shaw")):
        sd = synth.synthetic_state_dict(ptype, 0)
        model, cfg = rh.build_reference_model(yaml_name)
        model.load_state_dict(sd, strict=True)
        g = torch.Generator().manual_seed(123)
        z = torch.randn(2, 6, 128, 128, generator=g); x = torch.randn(2, 6, 256, 256, generator=g)
        from lib.utils.ce_utils import generate_mask_cond
        mask = generate_mask_cond(cfg, 2, z.device, None)
        with torch.no_grad():
            o = model(template=z, search=x, ce_template_mask=mask)
        out[name + "_mask"] = mask.numpy()
        for k in ("pred_boxes", "score_map", "size_map", "offset_map"):
            out[name + "_" + k] = o[k].numpy()
        out[name + "_feat_sample"] = o["backbone_feat"][:, ::16, ::32].numpy()
        out[name + "_feat_abs_sum"] = o["backbone_feat"].double().abs().sum().numpy()
        for i, r in enumerate(o["removed_indexes_s"]):
            out[name + "_removed_%d" % i] = r.numpy()
        out[name + "_attn_row27"] = o["attn"][:, :, 27, :].numpy()
    np.savez_compressed(os.path.join(OUT, "model_golden.npz"), **out)


def track_golden():
    out = {}
    for name, yaml_name, ptype, modality, seed, T in (
            ("deep_rgbd", "deep_rgbd", "vipt_deep", "rgbd", 11, 12),
            ("deep_rgbt", "deep_rgbt", "vipt_deep", "rgbt", 12, 6),
            ("shaw_rgbd", "shaw_rgbd", "vipt_shaw", "rgbd", 13, 6)):
        sd = synth.synthetic_state_dict(ptype, 0)
        tr = rh.make_tracker(yaml_name, sd)
        frames, box0, _ = synth.synthetic_sequence(seed, T, modality)
        tr.initialize(frames[0], {"init_bbox": list(box0)})
        boxes = [list(box0)]; scores = [1.0]
        for t in range(1, T):
            r = tr.track(frames[t])
            boxes.append(list(r["target_bbox"])); scores.append(r["best_score"])
        out[name + "_seed"] = np.int64(seed); out[name + "_T"] = np.int64(T)
        out[name + "_boxes"] = np.array(boxes, np.float64); out[name + "_scores"] = np.array(scores, np.float64)
        out[name + "_z_patch"] = tr.z_patch_arr
    np.savez_compressed(os.path.join(OUT, "track_golden.npz"), **out)


def peaked_golden():
    """Same two generators on the peaked-attention weight set; the tracking run is BASELINE.json configs[0]:
    one DepthTrack-shaped sequence, 50 tracked frames, batch 1 (sequence seed 300 = sequence 0 of
    oracle/parity_run.py)."""
    out = {}
    for name, yaml_name, ptype in (("deep", "deep_rgbt", "vipt_deep"), ("shaw", "shaw_rgbd", "vipt_shaw")):
        sd = synth.weight_set("peaked", ptype)
        model, cfg = rh.build_reference_model(yaml_name)
        model.load_state_dict(sd, strict=True)
        g = torch.Generator().manual_seed(321)
        z = torch.randn(4, 6, 128, 128, generator=g); x = torch.randn(4, 6, 256, 256, generator=g)
        from lib.utils.ce_utils import generate_mask_cond
        mask = generate_mask_cond(cfg, 4, z.device, None)
        with torch.no_grad():
            o = model(template=z, search=x, ce_template_mask=mask)
        for k in ("pred_boxes", "score_map", "size_map", "offset_map"):
            out[name + "_" + k] = o[k].numpy()
        out[name + "_feat_sample"] = o["backbone_feat"][:, ::16, ::32].numpy()
        for i, r in enumerate(o["removed_indexes_s"]):
            out[name + "_removed_%d" % i] = r.numpy()
        out[name + "_attn_row27"] = o["attn"][:, :, 27, :].numpy()
    T = 51
    sd = synth.weight_set("peaked", "vipt_deep")
    tr = rh.make_tracker("deep_rgbd", sd)
    frames, box0, _ = synth.synthetic_sequence(300, T, "rgbd")
    tr.initialize(frames[0], {"init_bbox": list(box0)})
    boxes = [list(box0)]; scores = [1.0]
    for t in range(1, T):
        r = tr.track(frames[t])
        boxes.append(list(r["target_bbox"])); scores.append(r["best_score"])
    out["track_seed"] = np.int64(300); out["track_T"] = np.int64(T)
    out["track_boxes"] = np.array(boxes, np.float64); out["track_scores"] = np.array(scores, np.float64)
    np.savez_compressed(os.path.join(OUT, "peaked_golden.npz"), **out)


def frames_golden():
    """get_x_frame (ViPT/lib/train/dataset/depth_utils.py:70-132) on small synthetic image files."""
    import tempfile
    import cv2
    # the unmodified file, loaded by path: the package __init__ pulls in dataset classes (pycocotools, lmdb ...)
    import importlib.util
    spec = importlib.util.spec_from_file_location("ref_depth_utils", os.path.join(rh.REF_ROOT, "lib", "train", "dataset",
                                                                                  "depth_utils.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    get_x_frame = mod.get_x_frame
    rng = np.random.default_rng(5)
    out = {}
    with tempfile.TemporaryDirectory() as d:
        for i, (H, W) in enumerate(((48, 64), (37, 53), (60, 41))):
            bgr = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
            xbgr = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
            yy, xx = np.mgrid[0:H, 0:W]
            depth = (900 + 40 * yy + 25 * xx + rng.integers(0, 300, (H, W))).astype(np.uint16)
            depth[rng.random((H, W)) < 0.05] = 0                    # holes, as in real depth maps
            depth[rng.random((H, W)) < 0.03] = 60000                # far outliers (exercise the clip)
            if i == 2:
                depth[:] = 1234                                     # constant image: max == min
            cv2.imwrite(os.path.join(d, "c.png"), bgr); cv2.imwrite(os.path.join(d, "x.png"), xbgr)
            cv2.imwrite(os.path.join(d, "d.png"), depth)
            out["bgr%d" % i], out["xbgr%d" % i], out["depth%d" % i] = bgr, xbgr, depth
            out["rgbrgb%d" % i] = get_x_frame(os.path.join(d, "c.png"), os.path.join(d, "x.png"), dtype="rgbrgb")
            for clip in (False, True):
                out["rgbcolormap%d_clip%d" % (i, int(clip))] = get_x_frame(
                    os.path.join(d, "c.png"), os.path.join(d, "d.png"), dtype="rgbcolormap", depth_clip=clip)
    np.savez_compressed(os.path.join(OUT, "frames_golden.npz"), **out)


if __name__ == "__main__":
    if "--frames-only" in sys.argv:
        frames_golden()
        sys.exit(0)
    torch.set_num_threads(8)
    if "--peaked-only" in sys.argv:
        peaked_golden()
        sys.exit(0)
    peaked_golden()
    crop_golden(); model_golden(); track_golden(); frames_golden()
    for f in ("crop_golden.npz", "model_golden.npz", "track_golden.npz"):
        print(f, os.path.getsize(os.path.join(OUT, f)))
