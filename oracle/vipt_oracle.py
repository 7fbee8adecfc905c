"""CPU fp32 restatement of the reference's ViPT tracking forward -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this module; the product path (the CUDA engine) never does and fails loudly without its .so.

Parity status: the reference has no tests / golden vectors for this path, so
this restatement is pinned against outputs of the reference itself, run unmodified in the build
container (oracle/ref_harness.py -> tests/golden/*.npz, generator tests/golden/make_golden.py).

A functional restatement over the reference's own `state_dict` (no nn.Module tree).  Follows
(paths relative to /root/reference/ViPT/):
  forward_features   lib/models/vipt/vit_ce_prompt.py:184-346
  Prompt_block/Fovea lib/models/vipt/vit_ce_prompt.py:22-71
  PatchEmbed         lib/models/layers/patch_embed.py:23-32
  Attention          lib/models/layers/attn.py:33-59
  CEBlock / CE       lib/models/layers/attn_blocks.py:9-104
  timm 0.5.4 Mlp     third-party (install_vipt.sh:63): fc1 -> GELU(erf) -> fc2
  forward_head       lib/models/vipt/ostrack_prompt.py:60-91
  CenterPredictor    lib/models/layers/head.py:130-201
  hann2d             lib/test/utils/hann.py:6-16
  ViPTTrack          lib/test/tracker/vipt.py:41-118
  clip_box           lib/utils/box_ops.py:97-106
  generate_mask_cond lib/utils/ce_utils.py:15-35
"""
import math

import numpy as np
import torch
import torch.nn.functional as F

from . import crop_oracle

NUM_HEADS = 12
LN_EPS = 1e-6
BN_EPS = 1e-5
CE_LOC = (3, 6, 9)
CE_KEEP = 0.7
FEAT_SZ = 16
SEARCH_SIZE = 256
TEMPLATE_SIZE = 128
SEARCH_FACTOR = 4.0
TEMPLATE_FACTOR = 2.0


# Optional operand rounding, used ONLY by the tests of the parity harness itself (tests/test_parity_harness.py): with
# OPERAND_ROUND = lambda t: t.half().float() the GEMM / attention operands are rounded the way a 16-bit tensor-core
# implementation rounds them, which gives a CPU stand-in for "an implementation with fp16-operand error".  None = the
# exact fp32 restatement (every parity number is taken against that).
OPERAND_ROUND = None


def _q(t):
    return t if OPERAND_ROUND is None else OPERAND_ROUND(t)


def _ln(x, sd, prefix):
    return F.layer_norm(x, (x.shape[-1],), sd[prefix + ".weight"], sd[prefix + ".bias"], LN_EPS)


def _patch_embed(img, sd, prefix):
    """patch_embed.py:28-30: conv k16 s16 -> (B, L, 768)."""
    y = F.conv2d(img, sd[prefix + ".proj.weight"], sd[prefix + ".proj.bias"], stride=16)
    return y.flatten(2).transpose(1, 2)


def _prompt_block(sd, i, a_tok, d_tok):
    """Prompt_block.forward on tokens (vit_ce_prompt.py:62-71); a_tok/d_tok: (B, L, 768) are the two
    channel halves.  Fovea softmax runs over the L positions per (batch, channel) (:40-44)."""
    p = "backbone.prompt_blocks.%d." % i
    w00 = sd[p + "conv0_0.weight"].flatten(1); w01 = sd[p + "conv0_1.weight"].flatten(1)
    w1 = sd[p + "conv1x1.weight"].flatten(1)
    x0 = F.linear(a_tok, w00, sd[p + "conv0_0.bias"])            # (B, L, 8)
    x1 = F.linear(d_tok, w01, sd[p + "conv0_1.bias"])
    mask = torch.softmax(x0 * sd[p + "fovea.smooth"], dim=1)     # over positions
    f = mask * x0 + x1
    return F.linear(f, w1, sd[p + "conv1x1.bias"])               # (B, L, 768)


def _attention(x, sd, prefix):
    """attn.py:33-59; returns (proj output, probs (B, H, N, N))."""
    B, N, C = x.shape
    qkv = _q(F.linear(_q(x), _q(sd[prefix + ".qkv.weight"]), sd[prefix + ".qkv.bias"]))
    qkv = qkv.reshape(B, N, 3, NUM_HEADS, C // NUM_HEADS).permute(2, 0, 3, 1, 4)
    q, k, v = qkv[0], qkv[1], qkv[2]
    attn = (q @ k.transpose(-2, -1)) * ((C // NUM_HEADS) ** -0.5)
    attn = attn.softmax(dim=-1)
    y = (_q(attn) @ v).transpose(1, 2).reshape(B, N, C)
    return _q(F.linear(_q(y), _q(sd[prefix + ".proj.weight"]), sd[prefix + ".proj.bias"])), attn


def _candidate_elimination(attn, tokens, lens_t, keep_ratio, global_index, box_mask_z, forced_removed=None):
    """attn_blocks.py:21-75.  `forced_removed` (B, n_removed) raster positions, test-only: the removal decision is
    taken from the caller instead of the top-k (parity protocol: the stages after a near-tie decision are compared
    against the reference arithmetic run on the implementation's kept set); order stays score-descending."""
    lens_s = attn.shape[-1] - lens_t
    bs, hn = attn.shape[:2]
    lens_keep = math.ceil(keep_ratio * lens_s)
    if lens_keep == lens_s:
        return tokens, global_index, None, None
    attn_t = attn[:, :, :lens_t, lens_t:]
    if box_mask_z is not None:
        m = box_mask_z.unsqueeze(1).unsqueeze(-1).expand(-1, hn, -1, lens_s)
        attn_t = attn_t[m].view(bs, hn, -1, lens_s)
    score = attn_t.mean(dim=2).mean(dim=1)                        # (B, L_s)
    sorted_attn, indices = torch.sort(score, dim=1, descending=True)
    if forced_removed is not None:
        gone = (global_index.gather(1, indices).unsqueeze(-1) == forced_removed.unsqueeze(1)).any(-1)   # sorted order
        assert int(gone.sum(1).min()) == int(gone.sum(1).max()) == lens_s - lens_keep
        order = torch.sort(gone.to(torch.int8), dim=1, stable=True).indices      # kept first, both score-descending
        indices = indices.gather(1, order)
    topk_idx, non_topk_idx = indices[:, :lens_keep], indices[:, lens_keep:]
    keep_index = global_index.gather(1, topk_idx)
    removed_index = global_index.gather(1, non_topk_idx)
    tokens_t, tokens_s = tokens[:, :lens_t], tokens[:, lens_t:]
    C = tokens.shape[-1]
    kept = tokens_s.gather(1, topk_idx.unsqueeze(-1).expand(bs, -1, C))
    return torch.cat([tokens_t, kept], dim=1), keep_index, removed_index, score


def _scatter_full(x_s, global_index_s, removed, lens_x):
    """vit_ce_prompt.py:276-285 / :325-334: kept tokens back to raster positions, zeros elsewhere."""
    B, _, C = x_s.shape
    if not removed:
        return x_s
    removed_cat = torch.cat(removed, dim=1)
    pad = torch.zeros(B, lens_x - x_s.shape[1], C, dtype=x_s.dtype, device=x_s.device)
    x_s = torch.cat([x_s, pad], dim=1)
    index_all = torch.cat([global_index_s, removed_cat], dim=1)
    return torch.zeros_like(x_s).scatter_(1, index_all.unsqueeze(-1).expand(B, -1, C), x_s)


def backbone_forward(sd, z_img, x_img, ce_template_mask, prompt_type="vipt_deep", trace=None, forced_removed=None):
    """VisionTransformerCE.forward_features.  z_img (B,6,128,128), x_img (B,6,256,256) fp32."""
    bb = "backbone."
    B = x_img.shape[0]
    z = _patch_embed(z_img[:, :3], sd, bb + "patch_embed")
    x = _patch_embed(x_img[:, :3], sd, bb + "patch_embed")
    z_dte = _patch_embed(z_img[:, 3:], sd, bb + "patch_embed_prompt")
    x_dte = _patch_embed(x_img[:, 3:], sd, bb + "patch_embed_prompt")
    # input prompt (:205-219): the same prompt_norms[0] on all four streams
    pn0 = bb + "prompt_norms.0"
    z_prompted = _prompt_block(sd, 0, _ln(z, sd, pn0), _ln(z_dte, sd, pn0))
    x_prompted = _prompt_block(sd, 0, _ln(x, sd, pn0), _ln(x_dte, sd, pn0))
    z = z + z_prompted + sd[bb + "pos_embed_z"]
    x = x + x_prompted + sd[bb + "pos_embed_x"]
    lens_z, lens_x = z.shape[1], x.shape[1]
    x = torch.cat([z, x], dim=1)                                   # cat_mode 'direct'
    gi_s = torch.arange(lens_x, dtype=torch.int64, device=x.device).repeat(B, 1)
    removed = []
    ce_scores = []
    ce_index = []          # raster position of every search slot entering each CE stage (ce_scores column order)
    attn = None
    for i in range(12):
        if i >= 1 and prompt_type == "vipt_deep":                  # :268-310
            x_ori = x
            full_s = _scatter_full(x[:, lens_z:], gi_s, removed, lens_x)
            full = torch.cat([x[:, :lens_z], full_s], dim=1)
            full = _ln(full, sd, bb + "prompt_norms.%d" % (i - 1))   # LN of a zero row = beta
            zp = _ln(z_prompted, sd, bb + "prompt_norms.%d" % i)
            xp = _ln(x_prompted, sd, bb + "prompt_norms.%d" % i)
            z_prompted = _prompt_block(sd, i, full[:, :lens_z], zp)
            x_prompted = _prompt_block(sd, i, full[:, lens_z:], xp)
            add_s = x_prompted.gather(1, gi_s.unsqueeze(-1).expand(B, -1, x.shape[-1]))
            x = x_ori + torch.cat([z_prompted, add_s], dim=1)
        blk = bb + "blocks.%d" % i
        if trace is not None:
            trace["x_in_%d" % i] = x.clone()
        y, attn = _attention(_ln(x, sd, blk + ".norm1"), sd, blk + ".attn")
        x = x + y
        if i in CE_LOC:
            ce_index.append(gi_s)
            x, gi_s, rem, score = _candidate_elimination(
                attn, x, lens_z, CE_KEEP, gi_s, ce_template_mask,
                None if forced_removed is None else forced_removed[len(removed)])     # entries may be None
            removed.append(rem)
            ce_scores.append(score)
        h = F.linear(_q(_ln(x, sd, blk + ".norm2")), _q(sd[blk + ".mlp.fc1.weight"]), sd[blk + ".mlp.fc1.bias"])
        h = _q(F.gelu(h))                                          # exact erf
        x = x + _q(F.linear(h, _q(sd[blk + ".mlp.fc2.weight"]), sd[blk + ".mlp.fc2.bias"]))
    x = _ln(x, sd, bb + "norm")
    full_s = _scatter_full(x[:, lens_z:], gi_s, removed, lens_x)     # exact zeros at removed positions
    feat = torch.cat([x[:, :lens_z], full_s], dim=1)
    aux = {"attn": attn, "removed_indexes_s": removed, "ce_scores": ce_scores, "ce_index": ce_index,
           "global_index_s": gi_s}
    return feat, aux


def _conv_bn_relu(x, sd, p):
    y = F.conv2d(x, sd[p + ".0.weight"], sd[p + ".0.bias"], padding=1)
    y = F.batch_norm(y, sd[p + ".1.running_mean"], sd[p + ".1.running_var"], sd[p + ".1.weight"],
                     sd[p + ".1.bias"], False, 0.0, BN_EPS)
    return F.relu(y)


def head_forward(sd, feat):
    """forward_head + CenterPredictor.get_score_map.  feat (B, 320, 768)."""
    B = feat.shape[0]
    opt = feat[:, -FEAT_SZ * FEAT_SZ:].transpose(1, 2).reshape(B, -1, FEAT_SZ, FEAT_SZ)
    outs = {}
    for tower in ("ctr", "offset", "size"):
        y = opt
        for li in range(1, 5):
            y = _conv_bn_relu(y, sd, "box_head.conv%d_%s" % (li, tower))
        outs[tower] = F.conv2d(y, sd["box_head.conv5_%s.weight" % tower], sd["box_head.conv5_%s.bias" % tower])
    sig = lambda t: torch.clamp(torch.sigmoid(t), min=1e-4, max=1 - 1e-4)
    return sig(outs["ctr"]), sig(outs["size"]), outs["offset"]


def cal_bbox(score_map, size_map, offset_map):
    """head.py:142-160 -> (bbox (B,4) normalised cx,cy,w,h ; max_score (B,1))."""
    max_score, idx = torch.max(score_map.flatten(1), dim=1, keepdim=True)
    idx_y = idx // FEAT_SZ
    idx_x = idx % FEAT_SZ
    idx2 = idx.unsqueeze(1).expand(idx.shape[0], 2, 1)
    size = size_map.flatten(2).gather(2, idx2)
    offset = offset_map.flatten(2).gather(2, idx2).squeeze(-1)
    bbox = torch.cat([(idx_x.to(torch.float) + offset[:, :1]) / FEAT_SZ,
                      (idx_y.to(torch.float) + offset[:, 1:]) / FEAT_SZ, size.squeeze(-1)], dim=1)
    return bbox, max_score


def model_forward(sd, template, search, ce_template_mask, prompt_type="vipt_deep", trace=None, forced_removed=None):
    """ViPTrack.forward (ostrack_prompt.py:39-58): same dict keys."""
    feat, aux = backbone_forward(sd, template, search, ce_template_mask, prompt_type, trace, forced_removed)
    score, size, offset = head_forward(sd, feat)
    bbox, _ = cal_bbox(score, size, offset)
    out = {"pred_boxes": bbox.view(-1, 1, 4), "score_map": score, "size_map": size, "offset_map": offset,
           "backbone_feat": feat}
    out.update(aux)
    return out


def hann2d(sz=FEAT_SZ):
    w = 0.5 * (1 - torch.cos((2 * math.pi / (sz + 1)) * torch.arange(1, sz + 1).float()))
    return w.reshape(1, 1, -1, 1) * w.reshape(1, 1, 1, -1)


def ce_template_mask(bs=1):
    """generate_mask_cond, CTR_POINT, 8x8 template grid: only token 3*8+3 = 27."""
    m = torch.zeros(bs, 8, 8)
    m[:, 3:4, 3:4] = 1
    return m.flatten(1).to(torch.bool)


def clip_box(box, H, W, margin=0):
    x1, y1, w, h = box
    x2, y2 = x1 + w, y1 + h
    x1 = min(max(0, x1), W - margin); x2 = min(max(margin, x2), W)
    y1 = min(max(0, y1), H - margin); y2 = min(max(margin, y2), H)
    return [x1, y1, max(margin, x2 - x1), max(margin, y2 - y1)]


class OracleTracker:
    """ViPTTrack (vipt.py:17-118) restated over a state_dict; same initialize/track contract."""

    def __init__(self, sd, prompt_type="vipt_deep", use_cv2=False):
        self.sd = {k: v.float() if v.dtype.is_floating_point else v for k, v in sd.items()}
        self.prompt_type = prompt_type
        self.window = hann2d()
        self.use_cv2 = use_cv2
        self.state = None
        self.last = None

    def initialize(self, image, info):
        patch, _ = crop_oracle.sample_target(image, info['init_bbox'], TEMPLATE_FACTOR, TEMPLATE_SIZE, self.use_cv2)
        self.z_patch = patch
        self.z_tensor = torch.from_numpy(crop_oracle.preprocess(patch))
        self.box_mask_z = ce_template_mask(1)
        self.state = list(info['init_bbox'])

    def track(self, image, info=None):
        H, W, _ = image.shape
        patch, rf = crop_oracle.sample_target(image, self.state, SEARCH_FACTOR, SEARCH_SIZE, self.use_cv2)
        search = torch.from_numpy(crop_oracle.preprocess(patch))
        with torch.no_grad():
            out = model_forward(self.sd, self.z_tensor, search, self.box_mask_z, self.prompt_type)
            response = self.window * out["score_map"]
            bbox, best = cal_bbox(response, out["size_map"], out["offset_map"])
        max_score = best[0][0].item()
        pred_box = (bbox.view(-1, 4).mean(dim=0) * SEARCH_SIZE / rf).tolist()
        cx_prev = self.state[0] + 0.5 * self.state[2]; cy_prev = self.state[1] + 0.5 * self.state[3]
        cx, cy, w, h = pred_box
        half = 0.5 * SEARCH_SIZE / rf
        cx_real = cx + (cx_prev - half); cy_real = cy + (cy_prev - half)
        self.state = clip_box([cx_real - 0.5 * w, cy_real - 0.5 * h, w, h], H, W, margin=10)
        self.last = {"out": out, "x_patch": patch, "resize_factor": rf}
        return {"target_bbox": self.state, "best_score": max_score}
