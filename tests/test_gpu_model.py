"""Model-level parity: vipt_b200 build_viptrack(...).forward vs the UNMODIFIED reference's outputs
(committed goldens, tests/golden/make_golden.py) and vs the CPU oracle on fresh seeded inputs.

Tolerances (BASELINE.json north_star; protocol in oracle/parity.py): CE kept-token index SETS bit-exact except at
documented near-ties (the reference's own scores of the disagreeing tokens within TAU_CE of its keep threshold -- then
the rest of the forward is checked against the reference arithmetic on the engine's kept set), score maps within 1e-2
of the map maximum, boxes (normalised) within 0.5 px of the 256-px search crop.  Every sample of every call is
checked; nothing is hand-picked."""
import os

import numpy as np
import pytest
import torch

from oracle import parity, vipt_oracle as vo
from vipt_b200 import synth
from vipt_b200.model import build_viptrack

pytestmark = pytest.mark.gpu
CFG = {"MODEL": {"BACKBONE": {"TYPE": "vit_base_patch16_224_ce_prompt", "STRIDE": 16, "CE_LOC": [3, 6, 9],
                              "CE_KEEP_RATIO": [0.7, 0.7, 0.7], "CE_TEMPLATE_RANGE": "CTR_POINT"},
                 "HEAD": {"TYPE": "CENTER", "NUM_CHANNELS": 256}},
       "DATA": {"SEARCH": {"SIZE": 256}, "TEMPLATE": {"SIZE": 128}}, "TRAIN": {"PROMPT": {"TYPE": "vipt_deep"}}}


def _model(ptype, weights="peaked"):
    cfg = dict(CFG, TRAIN={"PROMPT": {"TYPE": ptype}})
    m = build_viptrack(cfg, training=False)
    m.load_state_dict(synth.weight_set(weights, ptype), strict=True)
    return m.cuda().eval()


def _check_model(m, out, z, x, ptype, weights, golden=None, tol_map=1e-2):
    """Every sample of a model.forward call against the reference: CE sets equal (or differing only at documented
    near-ties, then everything downstream is checked against the reference arithmetic on the engine's kept sets),
    score / size / offset maps within 1e-2 of the map maximum, boxes within 0.5 px of the 256-px search crop."""
    B = z.shape[0]
    tau = parity.TAU_CE_BY_SET[weights]
    sd = synth.weight_set(weights, ptype)
    mask = vo.ce_template_mask(B)
    with torch.no_grad():
        ref = vo.model_forward(sd, z, x, mask, ptype)
    if golden is not None:      # the pinned oracle equals the unmodified reference on these very inputs
        for k in ("score_map", "size_map", "offset_map", "pred_boxes"):
            np.testing.assert_allclose(ref[k].numpy(), golden[k], rtol=0, atol=5e-5)
    eng = m.engine(B)
    removed = np.zeros((3, B, 256), np.int64)
    for s in range(3):
        r = out["removed_indexes_s"][s].cpu().numpy()
        assert r.shape == (B, parity.N_S[s] - parity.KEEP[s])
        removed[s, :, :r.shape[1]] = r
    ce_score = eng.buffer("ce_score").cpu().numpy()
    gidx = parity.gpu_stage_index(ce_score, removed)
    status = [parity._ce_compare(ref, removed, ce_score, gidx, b, tau) for b in range(B)]
    assert all(st[0] != "mismatch" for st in status), [st[0] for st in status]
    redo = [b for b in range(B) if status[b][0] == "near_tie"]
    forced = None
    if redo:
        fr = [torch.from_numpy(removed[s, redo, :parity.N_S[s] - parity.KEEP[s]]) for s in range(3)]
        with torch.no_grad():
            forced = vo.model_forward(sd, z[redo], x[redo], mask[redo], ptype, forced_removed=fr)
    n_direct = 0
    for b in range(B):
        r, k = (ref, b) if status[b][0] == "equal" else (forced, redo.index(b))
        n_direct += status[b][0] == "equal"
        for name in ("score_map", "size_map", "offset_map"):
            err = (out[name].cpu()[b] - r[name][k]).abs().max().item() / r[name][k].abs().max().item()
            assert err < tol_map, (b, name, err, status[b][0])
        sm = r["score_map"][k].flatten()
        top = sm.topk(2).values
        if (top[0] - top[1]) / top[0] > parity.TAU_AM:
            assert out["score_map"].cpu()[b].flatten().argmax() == sm.argmax()
            box_px = (out["pred_boxes"].cpu().view(B, 4)[b] - r["pred_boxes"].view(-1, 4)[k]).abs().max().item() * 256
            assert box_px < 0.5, (b, box_px)
    return n_direct, ref


@pytest.mark.parametrize("name,ptype", [("deep", "vipt_deep"), ("shaw", "vipt_shaw")])
def test_forward_matches_reference_golden(golden_dir, name, ptype):
    """build_viptrack(...).forward on the inputs of tests/golden/peaked_golden.npz (outputs of the unmodified
    reference, peaked-attention weight set): output dict, shapes, and every sample within the parity bar."""
    g = np.load(os.path.join(golden_dir, "peaked_golden.npz"))
    gen = torch.Generator().manual_seed(321)
    z = torch.randn(4, 6, 128, 128, generator=gen); x = torch.randn(4, 6, 256, 256, generator=gen)
    m = _model(ptype, "peaked")
    out = m.forward(template=z.cuda(), search=x.cuda(), ce_template_mask=vo.ce_template_mask(4).cuda(),
                    return_backbone_feat=True)
    assert set(out) >= {"pred_boxes", "score_map", "size_map", "offset_map", "removed_indexes_s", "backbone_feat"}
    assert out["pred_boxes"].shape == (4, 1, 4) and out["score_map"].shape == (4, 1, 16, 16)
    assert out["size_map"].shape == (4, 2, 16, 16) and out["offset_map"].shape == (4, 2, 16, 16)
    assert [tuple(r.shape) for r in out["removed_indexes_s"]] == [(4, 76), (4, 54), (4, 37)]
    gold = {k: g[name + "_" + k] for k in ("score_map", "size_map", "offset_map", "pred_boxes")}
    n_direct, ref = _check_model(m, out, z, x, ptype, "peaked", gold)
    # backbone features of the samples whose kept sets equal the reference's: kept rows close, removed rows EXACT zeros
    feat = out["backbone_feat"].cpu()
    for b in range(4):
        if all(set(out["removed_indexes_s"][s][b].tolist()) == set(ref["removed_indexes_s"][s][b].tolist()) for s in range(3)):
            np.testing.assert_allclose(feat[b, ::16, ::32].numpy(), g[name + "_feat_sample"][b], atol=3e-2)
        gone = torch.cat([out["removed_indexes_s"][s][b].cpu() for s in range(3)]) + 64
        assert feat[b, gone].abs().max().item() == 0.0


@pytest.mark.parametrize("weights", ["peaked", "flat"])
def test_forward_batch_vs_oracle(weights):
    """fresh inputs, batch 5 (not a tile multiple): every sample against the pinned CPU oracle."""
    B = 5
    gen = torch.Generator().manual_seed(7)
    z = torch.randn(B, 6, 128, 128, generator=gen); x = torch.randn(B, 6, 256, 256, generator=gen)
    m = _model("vipt_deep", weights)
    out = m.forward(z.cuda(), x.cuda(), vo.ce_template_mask(B).cuda())
    _check_model(m, out, z, x, "vipt_deep", weights)


def test_api_errors():
    m = build_viptrack(CFG)
    with pytest.raises(RuntimeError):
        m.load_state_dict({"backbone.norm.weight": torch.zeros(768)}, strict=True)
    with pytest.raises(ValueError):
        build_viptrack({"MODEL": {"HEAD": {"TYPE": "MLPX"}}})


def test_outlier_channels_fp16_range():
    """16-bit operand range: weights with outlier LayerNorm gains (x40 on a few channels of four
    blocks, fc1 columns x4 -- the pattern trained ViTs show) drive the fp16 GEMM operands two orders of magnitude above
    the random-init range.  The saturating converts (cvt.rn.satfinite.f16x2.f32) must stay far from their clamp in
    every fp16 buffer, and the forward must still meet the parity bar on every sample."""
    B = 4
    gen = torch.Generator().manual_seed(11)
    z = torch.randn(B, 6, 128, 128, generator=gen); x = torch.randn(B, 6, 256, 256, generator=gen)
    m = _model("vipt_deep", "outlier")
    out = m.forward(z.cuda(), x.cuda(), vo.ce_template_mask(B).cuda())
    eng = m.engine(B)
    peak = {}
    for name in ("xn", "qkv", "att", "hbuf", "dbuf", "feat16", "h1"):
        t = eng.buffer(name).float()
        assert torch.isfinite(t).all(), name
        peak[name] = t.abs().max().item()
    print("fp16 buffer maxima:", {k: round(v, 1) for k, v in peak.items()})
    assert peak["xn"] > 50.0                      # the outliers are really there ...
    assert max(peak.values()) < 0.25 * 65504      # ... and two binades below the fp16 clamp
    # Rounding errors of 16-bit operands grow with the outliers: on these very inputs the fp32 reference itself, with
    # its GEMM / attention operands rounded to fp16 (vipt_oracle.OPERAND_ROUND), moves its score maps by up to 7.5e-3
    # of the maximum (3e-3 .. 7.5e-3 over the four samples).  The stress case therefore gets 2e-2, the regular weight
    # sets keep the north-star 1e-2.
    _check_model(m, out, z, x, "vipt_deep", "outlier", tol_map=2e-2)


@pytest.mark.parametrize("weights,B,folds", [("outlier", 48, (9, 11)), ("peaked", 48, (9, 11)), ("peaked", 24, (3, 3))])
def test_large_batch_folds_vs_oracle(weights, B, folds):
    """The large-batch arithmetic (LayerNorm fold: fp16(x) instead of fp16(LN(x)) as GEMM operand; prompt fold: row sums in
    the fc2 epilogue, rank-8 update as an extra k-block) at the model boundary: 24 samples -- enough rows for the 2-CTA
    GEMMs, so every fold is active -- against the pinned CPU oracle, every sample, also with the outlier-channel weights
    (residual-stream values ~50: the un-normalised fp16 operand must keep the parity bar of the stress case).  24 samples:
    only the 320-token blocks have enough rows, folded and separate-kernel blocks alternate inside one forward."""
    gen = torch.Generator().manual_seed(23)
    z = torch.randn(B, 6, 128, 128, generator=gen); x = torch.randn(B, 6, 256, 256, generator=gen)
    m = _model("vipt_deep", weights)
    out = m.forward(z.cuda(), x.cuda(), vo.ce_template_mask(B).cuda())
    eng = m.engine(B)
    assert (eng.info("lnfold_layers"), eng.info("promptfold_layers")) == folds
    for name in ("xn", "qkv", "att", "hbuf", "feat16", "h1"):
        t = eng.buffer(name).float()
        assert torch.isfinite(t).all(), name
        assert t.abs().max().item() < 0.25 * 65504, name
    # Outlier stress set: over these 48 samples the SEPARATE-kernel arithmetic (VIPT_NO_LNFOLD=1) reaches 2.6e-2 of the map
    # maximum (and one CE decision outside the tie band), the folded arithmetic 2.4e-2 with none -- measured with
    # scripts/fold_error_ab.py on the same inputs; the 4-sample stress test above sees 7.5e-3.  The bar for
    # this set at this sample count is therefore 3e-2; the regular weight set keeps the north-star 1e-2 (measured 3.2e-3).
    _check_model(m, out, z, x, "vipt_deep", weights, tol_map=3e-2 if weights == "outlier" else 1e-2)


@pytest.mark.parametrize("switch,want", [("VIPT_NO_LNFOLD", (0, 0)), ("VIPT_NO_PROMPT_FOLD", (9, 0)), ("VIPT_PROMPT_FOLD_NO_CE", (9, 8))])
def test_fold_switches(monkeypatch, switch, want):
    """LayerNorm fold / prompt fold (csrc/gemm2.cu, prompt_mid.cu) against the same engine with the fold switched off
    (environment read once at engine creation): the states both forms materialise before the first candidate
    elimination -- qkv of block 1, the residual stream after attn.proj + norm2 of block 2, fc1's output of block 2 --
    agree to fp16-operand accuracy at the large-batch size where the folds are active."""
    from vipt_b200.engine import NativeEngine
    B = 64
    sd = synth.weight_set("peaked", "vipt_deep")
    g = torch.Generator().manual_seed(11)
    z = torch.randn(B, 6, 128, 128, generator=g).cuda()
    x = torch.randn(B, 6, 256, 256, generator=g).cuda()
    on = NativeEngine(sd, B, "vipt_deep")
    assert (on.info("lnfold_layers"), on.info("promptfold_layers")) == (9, 11)
    monkeypatch.setenv(switch, "1")
    off = NativeEngine(sd, B, "vipt_deep")
    monkeypatch.delenv(switch)
    assert (off.info("lnfold_layers"), off.info("promptfold_layers")) == want
    width = {"qkv": 2304, "hbuf": 3072, "xA": 768}
    for stage, name in ((4 + 7, "qkv"), (7 + 14, "xA"), (8 + 14, "hbuf")):
        on.forward(z, x, debug_stop=stage)
        off.forward(z, x, debug_stop=stage)
        n = B * 320 * width[name]
        a, b = on.buffer(name).flatten()[:n].float(), off.buffer(name).flatten()[:n].float()
        rel = ((a - b).abs().max() / b.abs().max()).item()
        assert rel < 5e-3, (switch, name, rel)


@pytest.mark.parametrize("B", [2, 48])
def test_prompt_stage_against_oracle(B):
    """Per-stage check of the MCP prompt blocks (vit_ce_prompt.py:205-219 input prompt, :268-310 deep stage, Prompt_block /
    Fovea :22-71) against the pinned CPU oracle, not only through the final maps: the oracle's residual stream entering
    blocks 1 and 2 (after the input prompt resp. the first deep-prompt stage) gives qkv = norm1(x) Wqkv^T + b of that
    block, which is what the engine materialises in both forms -- batch 2: prompt_coop_kernel + the LayerNorm it writes;
    batch 48: prompt fold (fc2 epilogue sums, prompt_mid_kernel, rank-8 update as the extra k-block of the qkv GEMM)."""
    from vipt_b200.engine import NativeEngine
    sd = synth.weight_set("peaked", "vipt_deep")
    gen = torch.Generator().manual_seed(31)
    z = torch.randn(B, 6, 128, 128, generator=gen); x = torch.randn(B, 6, 256, 256, generator=gen)
    trace = {}
    with torch.no_grad():
        vo.backbone_forward(sd, z, x, vo.ce_template_mask(B), "vipt_deep", trace=trace)
    eng = NativeEngine(sd, B, "vipt_deep")
    assert eng.info("promptfold_layers") == (11 if B == 48 else 0)
    for i in (0, 1, 2):
        p = "backbone.blocks.%d." % i
        xin = trace["x_in_%d" % i]                                        # (B, 320, 768) entering block i
        ln = torch.nn.functional.layer_norm(xin, (768,), sd[p + "norm1.weight"], sd[p + "norm1.bias"], 1e-6)
        ref = ln @ sd[p + "attn.qkv.weight"].t() + sd[p + "attn.qkv.bias"]
        eng.forward(z.cuda(), x.cuda(), debug_stop=4 + 7 * i)
        got = eng.buffer("qkv").flatten()[:B * 320 * 2304].view(B, 320, 2304).float().cpu()
        err = ((got - ref).abs().max() / ref.abs().max()).item()
        assert err < 5e-3, (B, i, err)
