"""Drop-in for the reference's model API: `build_viptrack(cfg, training=False)` and
`ViPTrack.forward(template, search, ce_template_mask)` (ViPT/lib/models/vipt/ostrack_prompt.py:39-58,94-138).

Same call signature, same output dict keys, same `load_state_dict(sd, strict=True)` key set, and the
`box_head.cal_bbox(score, size, offset, return_score)` attribute that ViPTTrack calls from outside
(ViPT/lib/test/tracker/vipt.py:81).  The computation runs in libvipt_b200.so; there is no CPU path.

Documented deviations: `attn` (last-layer probabilities, never materialised by the fused
attention kernel) is not returned; `backbone_feat` is returned only with `return_backbone_feat=True`.
"""
import numpy as np
import torch

from . import synth
from .engine import NativeEngine

CE_TOKEN = 27   # generate_mask_cond, CTR_POINT, 8x8 template grid (ViPT/lib/utils/ce_utils.py:22-35)


def _cfg_get(cfg, path, default=None):
    cur = cfg
    for p in path.split("."):
        if cur is None:
            return default
        cur = cur.get(p) if isinstance(cur, dict) else getattr(cur, p, None)
    return default if cur is None else cur


class _BoxHead:
    """The part of CenterPredictor that callers touch directly (head.py:142-160)."""
    feat_sz = 16

    def cal_bbox(self, score_map_ctr, size_map, offset_map, return_score=False):
        max_score, idx = torch.max(score_map_ctr.flatten(1), dim=1, keepdim=True)
        idx_y = idx // self.feat_sz
        idx_x = idx % self.feat_sz
        idx = idx.unsqueeze(1).expand(idx.shape[0], 2, 1)
        size = size_map.flatten(2).gather(dim=2, index=idx)
        offset = offset_map.flatten(2).gather(dim=2, index=idx).squeeze(-1)
        bbox = torch.cat([(idx_x.to(torch.float) + offset[:, :1]) / self.feat_sz,
                          (idx_y.to(torch.float) + offset[:, 1:]) / self.feat_sz,
                          size.squeeze(-1)], dim=1)
        if return_score:
            return bbox, max_score
        return bbox


class ViPTrackB200:
    """Mirror of ViPTrack (ostrack_prompt.py:17) over the native engine."""

    def __init__(self, prompt_type="vipt_deep"):
        self.prompt_type = prompt_type
        self.box_head = _BoxHead()
        self._sd = None
        self._engines = {}
        self.training = False

    # nn.Module-like surface used by the reference callers (vipt.py:20-24)
    def load_state_dict(self, state_dict, strict=True):
        want = set(synth.synthetic_state_dict_keys(self.prompt_type))
        have = set(state_dict.keys())
        if strict and want != have:
            raise RuntimeError("Error(s) in loading state_dict for ViPTrack: missing %s unexpected %s"
                               % (sorted(want - have)[:5], sorted(have - want)[:5]))
        self._sd = {k: v.detach().cpu() for k, v in state_dict.items()}
        self._engines = {}

    def cuda(self, device=None):
        return self

    def eval(self):
        return self

    def engine(self, batch):
        if self._sd is None:
            raise RuntimeError("load_state_dict() first")
        if batch not in self._engines:
            self._engines[batch] = NativeEngine(self._sd, batch, self.prompt_type)
        return self._engines[batch]

    def forward(self, template, search, ce_template_mask=None, ce_keep_rate=None, return_last_attn=False,
                return_backbone_feat=False):
        if ce_keep_rate is not None:
            raise NotImplementedError("ce_keep_rate override is training-only (ViPT/lib/train/actors/vipt.py:78)")
        B = search.shape[0]
        if ce_template_mask is not None:
            m = ce_template_mask.detach().cpu().numpy().reshape(B, -1)
            ok = m.shape[1] == 64 and (m.sum(1) == 1).all() and m[:, CE_TOKEN].all()
            if not ok:
                raise NotImplementedError("only CE_TEMPLATE_RANGE = CTR_POINT (token 27) is supported")
        else:
            raise NotImplementedError("CE_TEMPLATE_RANGE = ALL (mask None) is not supported; every shipped yaml "
                                      "uses CTR_POINT")
        eng = self.engine(B)
        eng.forward(template, search, want_feat=return_backbone_feat)
        n_removed = [76, 54, 37]
        removed = eng.buffer("removed")
        out = {
            "pred_boxes": eng.buffer("pred_boxes").clone().view(B, 1, 4),
            "score_map": eng.buffer("score_map").clone(),
            "size_map": eng.buffer("size_map").clone(),
            "offset_map": eng.buffer("offset_map").clone(),
            "removed_indexes_s": [removed[i, :, :n_removed[i]].to(torch.int64) for i in range(3)],
        }
        if return_backbone_feat:
            out["backbone_feat"] = eng.buffer("feat32").clone()
        return out

    __call__ = forward


def build_viptrack(cfg, training=False):
    """Same factory signature as ostrack_prompt.py:94.  Only the configuration every shipped yaml uses is
    supported: vit_base_patch16_224_ce_prompt, CENTER head, 128/256, CE [3,6,9] x 0.7, CTR_POINT."""
    if training:
        raise NotImplementedError("inference only")
    btype = _cfg_get(cfg, "MODEL.BACKBONE.TYPE", "vit_base_patch16_224_ce_prompt")
    if btype != "vit_base_patch16_224_ce_prompt":
        raise NotImplementedError(btype)
    if _cfg_get(cfg, "MODEL.HEAD.TYPE", "CENTER") != "CENTER":
        raise ValueError("HEAD TYPE %s is not supported." % _cfg_get(cfg, "MODEL.HEAD.TYPE"))
    checks = (("MODEL.BACKBONE.STRIDE", 16), ("MODEL.HEAD.NUM_CHANNELS", 256), ("DATA.SEARCH.SIZE", 256),
              ("DATA.TEMPLATE.SIZE", 128))
    for path, want in checks:
        got = _cfg_get(cfg, path, want)
        if got != want:
            raise NotImplementedError("%s = %r (engine is built for %r)" % (path, got, want))
    if list(_cfg_get(cfg, "MODEL.BACKBONE.CE_LOC", [3, 6, 9])) != [3, 6, 9] or \
            [float(v) for v in _cfg_get(cfg, "MODEL.BACKBONE.CE_KEEP_RATIO", [0.7] * 3)] != [0.7] * 3:
        raise NotImplementedError("CE_LOC / CE_KEEP_RATIO other than [3,6,9] x 0.7")
    return ViPTrackB200(_cfg_get(cfg, "TRAIN.PROMPT.TYPE", "vipt_deep"))
