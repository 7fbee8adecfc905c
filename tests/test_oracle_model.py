"""Pins oracle/vipt_oracle.py against golden vectors produced by the UNMODIFIED reference
(tests/golden/make_golden.py), and against the live reference when /root/reference exists."""
import os

import numpy as np
import pytest
import torch

from oracle import ref_harness, vipt_oracle as vo
from vipt_b200 import synth


@pytest.mark.parametrize("name,ptype", [("deep", "vipt_deep"), ("shaw", "vipt_shaw")])
def test_model_forward_matches_reference_golden(golden_dir, name, ptype):
    g = np.load(os.path.join(golden_dir, "model_golden.npz"))
    sd = synth.synthetic_state_dict(ptype, 0)
    gen = torch.Generator().manual_seed(123)
    z = torch.randn(2, 6, 128, 128, generator=gen); x = torch.randn(2, 6, 256, 256, generator=gen)
    mask = vo.ce_template_mask(2)
    assert np.array_equal(mask.numpy(), g[name + "_mask"])
    with torch.no_grad():
        o = vo.model_forward(sd, z, x, mask, ptype)
    for k in ("pred_boxes", "score_map", "size_map", "offset_map"):
        np.testing.assert_allclose(o[k].numpy(), g[name + "_" + k], rtol=0, atol=2e-5)
    for i in range(3):   # index work: bit-exact, order included
        assert np.array_equal(o["removed_indexes_s"][i].numpy(), g[name + "_removed_%d" % i])
    np.testing.assert_allclose(o["backbone_feat"][:, ::16, ::32].numpy(), g[name + "_feat_sample"], atol=1e-4)
    np.testing.assert_allclose(o["attn"][:, :, 27, :].numpy(), g[name + "_attn_row27"], atol=1e-6)


@pytest.mark.parametrize("name,ptype,modality", [("deep_rgbd", "vipt_deep", "rgbd"), ("deep_rgbt", "vipt_deep", "rgbt"),
                                                 ("shaw_rgbd", "vipt_shaw", "rgbd")])
def test_tracker_matches_reference_golden(golden_dir, name, ptype, modality):
    g = np.load(os.path.join(golden_dir, "track_golden.npz"))
    T = min(int(g[name + "_T"]), 5)
    frames, box0, _ = synth.synthetic_sequence(int(g[name + "_seed"]), int(g[name + "_T"]), modality)
    tr = vo.OracleTracker(synth.synthetic_state_dict(ptype, 0), ptype)
    tr.initialize(frames[0], {"init_bbox": list(box0)})
    assert np.array_equal(tr.z_patch, g[name + "_z_patch"])
    for t in range(1, T):
        r = tr.track(frames[t])
        np.testing.assert_allclose(r["target_bbox"], g[name + "_boxes"][t], atol=5e-3)   # free-running, fp32 noise
        assert abs(r["best_score"] - g[name + "_scores"][t]) < 1e-4


@pytest.mark.skipif(not ref_harness.available(), reason="reference checkout not present (GPU box)")
def test_state_dict_keys_match_reference():
    for yaml_name, ptype, n in (("deep_rgbe", "vipt_deep", 352), ("shaw_rgbt", "vipt_shaw", 253)):
        model, _ = ref_harness.build_reference_model(yaml_name)
        sd = synth.synthetic_state_dict(ptype, 0)
        assert len(sd) == n
        model.load_state_dict(sd, strict=True)
