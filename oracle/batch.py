"""Batched form of the oracle tracker -- TEST INFRASTRUCTURE ONLY (tests/, scripts/parity_report.py, bench.py CPU legs).

`OracleBatchTracker` advances n independent sequences in lock-step exactly like n instances of
`vipt_oracle.OracleTracker` (ViPT/lib/test/tracker/vipt.py:41-118): per-sequence crop (sample_target +
PreprocessorMM), ONE batched fp32 forward, per-sequence box decode / map_box_back / clip_box.  Batching the forward
only changes how many rows the CPU GEMMs see; every sequence's arithmetic is the reference's.
"""
import numpy as np
import torch

from . import crop_oracle as co
from . import vipt_oracle as vo


def decode_box(bbox_norm, state, rf, H, W):
    """vipt.py:80-88 + map_box_back (:112-118) + clip_box: normalised (cx, cy, w, h) -> image [x, y, w, h]."""
    cx, cy, w, h = (np.asarray(bbox_norm, np.float64) * vo.SEARCH_SIZE / rf).tolist()
    half = 0.5 * vo.SEARCH_SIZE / rf
    cxr = cx + (state[0] + 0.5 * state[2] - half)
    cyr = cy + (state[1] + 0.5 * state[3] - half)
    return vo.clip_box([cxr - 0.5 * w, cyr - 0.5 * h, w, h], H, W, margin=10)


class OracleBatchTracker:
    def __init__(self, sd, prompt_type="vipt_deep", use_cv2=False):
        self.sd = {k: v.float() if v.dtype.is_floating_point else v for k, v in sd.items()}
        self.prompt_type = prompt_type
        self.use_cv2 = use_cv2
        self.window = vo.hann2d()
        self.state = None
        self.last = None

    def initialize(self, images, boxes):
        self.n = len(images)
        z = [co.preprocess(co.sample_target(im, list(b), vo.TEMPLATE_FACTOR, vo.TEMPLATE_SIZE, self.use_cv2)[0])
             for im, b in zip(images, boxes)]
        self.z = torch.from_numpy(np.concatenate(z, 0))
        self.mask = vo.ce_template_mask(self.n)
        self.state = [list(map(float, b)) for b in boxes]
        self.hw = [im.shape[:2] for im in images]

    def crops(self, images):
        crops, rfs = [], []
        for i, im in enumerate(images):
            p, rf = co.sample_target(im, self.state[i], vo.SEARCH_FACTOR, vo.SEARCH_SIZE, self.use_cv2)
            crops.append(co.preprocess(p)); rfs.append(rf)
        return torch.from_numpy(np.concatenate(crops, 0)), rfs

    def forward(self, x, forced_removed=None, rows=None):
        """batched model forward on search crops x; rows = subset of sequences (list of indices) or None."""
        z, mask = self.z, self.mask
        if rows is not None:
            z, mask, x = z[rows], mask[rows], x[rows]
        with torch.no_grad():
            out = vo.model_forward(self.sd, z, x, mask, self.prompt_type, forced_removed=forced_removed)
            out["response"] = self.window * out["score_map"]
            bbox, best = vo.cal_bbox(out["response"], out["size_map"], out["offset_map"])
        out["bbox"] = bbox
        out["best"] = best[:, 0]
        return out

    def boxes_from(self, out, rfs, rows=None):
        rows = list(range(self.n)) if rows is None else rows
        return np.array([decode_box(out["bbox"][k].numpy(), self.state[i], rfs[i], self.hw[i][0], self.hw[i][1])
                         for k, i in enumerate(rows)], np.float64)

    def track(self, images):
        """-> (boxes (n,4) f64, scores (n,) f32); self.last keeps the forward outputs, crops and resize factors."""
        x, rfs = self.crops(images)
        out = self.forward(x)
        prev = [list(s) for s in self.state]
        boxes = self.boxes_from(out, rfs)
        self.state = [list(b) for b in boxes]
        self.last = {"out": out, "x": x, "rfs": rfs, "prev_state": prev}
        return boxes, out["best"].numpy().copy()

    def set_state(self, boxes):
        self.state = [list(map(float, b)) for b in boxes]
