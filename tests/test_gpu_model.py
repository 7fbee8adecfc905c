"""Model-level parity: vipt_b200 build_viptrack(...).forward vs the UNMODIFIED reference's outputs
(committed goldens, tests/golden/make_golden.py) and vs the CPU oracle on fresh seeded inputs.

Tolerances (BASELINE.json north_star): CE kept-token index SETS bit-exact except at documented
near-ties (reference score gap at the keep boundary < TAU_CE relative), score maps within 1e-2 of the
map maximum, boxes (normalised) within 0.5 px of the 256-px search crop."""
import os

import numpy as np
import pytest
import torch

from oracle import near_tie, vipt_oracle as vo
from vipt_b200 import synth
from vipt_b200.model import build_viptrack

pytestmark = pytest.mark.gpu
CFG = {"MODEL": {"BACKBONE": {"TYPE": "vit_base_patch16_224_ce_prompt", "STRIDE": 16, "CE_LOC": [3, 6, 9],
                              "CE_KEEP_RATIO": [0.7, 0.7, 0.7], "CE_TEMPLATE_RANGE": "CTR_POINT"},
                 "HEAD": {"TYPE": "CENTER", "NUM_CHANNELS": 256}},
       "DATA": {"SEARCH": {"SIZE": 256}, "TEMPLATE": {"SIZE": 128}}, "TRAIN": {"PROMPT": {"TYPE": "vipt_deep"}}}


def _model(ptype):
    cfg = dict(CFG, TRAIN={"PROMPT": {"TYPE": ptype}})
    m = build_viptrack(cfg, training=False)
    m.load_state_dict(synth.synthetic_state_dict(ptype, 0), strict=True)
    return m.cuda().eval()


def ce_status(out, ref_removed, ref_scores):
    st = near_tie.ce_set_status([r.cpu() for r in out["removed_indexes_s"]], ref_removed, ref_scores)
    assert "mismatch" not in st, st          # CE kept sets: bit-exact except at documented near-ties
    return st


@pytest.mark.parametrize("name,ptype", [("deep", "vipt_deep"), ("shaw", "vipt_shaw")])
def test_forward_matches_reference_golden(golden_dir, name, ptype):
    g = np.load(os.path.join(golden_dir, "model_golden.npz"))
    gen = torch.Generator().manual_seed(123)
    z = torch.randn(2, 6, 128, 128, generator=gen); x = torch.randn(2, 6, 256, 256, generator=gen)
    m = _model(ptype)
    out = m.forward(template=z.cuda(), search=x.cuda(), ce_template_mask=torch.from_numpy(g[name + "_mask"]).cuda(),
                    return_backbone_feat=True)
    assert set(out) >= {"pred_boxes", "score_map", "size_map", "offset_map", "removed_indexes_s", "backbone_feat"}
    assert out["pred_boxes"].shape == (2, 1, 4) and out["score_map"].shape == (2, 1, 16, 16)
    assert out["size_map"].shape == (2, 2, 16, 16) and out["offset_map"].shape == (2, 2, 16, 16)
    assert [tuple(r.shape) for r in out["removed_indexes_s"]] == [(2, 76), (2, 54), (2, 37)]
    with torch.no_grad():   # the pinned oracle supplies the reference CE scores for the near-tie criterion
        orc = vo.model_forward(synth.synthetic_state_dict(ptype, 0), z, x, vo.ce_template_mask(2), ptype)
    st = ce_status(out, [torch.from_numpy(g[name + "_removed_%d" % i]) for i in range(3)], orc["ce_scores"])
    eq = [b for b in range(2) if st[b] == "equal"]
    assert eq, "every sequence of the golden sits at a CE near-tie: %s" % st
    for k in ("score_map", "size_map", "offset_map"):
        ref = torch.from_numpy(g[name + "_" + k])[eq]
        err = (out[k].cpu()[eq] - ref).abs().max().item() / ref.abs().max().item()
        assert err < 1e-2, (k, err)
    ref_sm = torch.from_numpy(g[name + "_score_map"])
    for b in eq:
        if near_tie.argmax_gap(ref_sm[b:b + 1]).item() > near_tie.TAU_AM:
            assert out["score_map"].cpu()[b].flatten().argmax() == ref_sm[b].flatten().argmax()
            box_px = (out["pred_boxes"].cpu()[b] - torch.from_numpy(g[name + "_pred_boxes"])[b]).abs().max().item() * 256
            assert box_px < 0.5, box_px
    feat = out["backbone_feat"].cpu()[:, ::16, ::32]
    np.testing.assert_allclose(feat.numpy()[eq], g[name + "_feat_sample"][eq], atol=2e-2)


def test_forward_batch_vs_oracle():
    """fresh inputs, batch 5 (not a tile multiple): against the pinned CPU oracle."""
    B = 5
    gen = torch.Generator().manual_seed(7)
    z = torch.randn(B, 6, 128, 128, generator=gen); x = torch.randn(B, 6, 256, 256, generator=gen)
    sd = synth.synthetic_state_dict("vipt_deep", 0)
    mask = vo.ce_template_mask(B)
    with torch.no_grad():
        ref = vo.model_forward(sd, z, x, mask, "vipt_deep")
    out = _model("vipt_deep").forward(z.cuda(), x.cuda(), mask.cuda())
    st = ce_status(out, ref["removed_indexes_s"], ref["ce_scores"])
    eq = [b for b in range(B) if st[b] == "equal"]
    assert len(eq) >= 2, st
    err = (out["score_map"].cpu()[eq] - ref["score_map"][eq]).abs().max().item() / ref["score_map"][eq].max().item()
    assert err < 1e-2, err
    for b in eq:
        if near_tie.argmax_gap(ref["score_map"][b:b + 1]).item() > near_tie.TAU_AM:
            assert (out["pred_boxes"].cpu().view(B, 4)[b] - ref["pred_boxes"].view(B, 4)[b]).abs().max().item() * 256 < 0.5


def test_api_errors():
    m = build_viptrack(CFG)
    with pytest.raises(RuntimeError):
        m.load_state_dict({"backbone.norm.weight": torch.zeros(768)}, strict=True)
    with pytest.raises(ValueError):
        build_viptrack({"MODEL": {"HEAD": {"TYPE": "MLPX"}}})
