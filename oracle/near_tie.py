"""Near-tie bookkeeping for the parity protocol -- TEST INFRASTRUCTURE ONLY.

A frame is a *documented near-tie* when the fp32 reference itself is within TAU of making a different
discrete decision, so an fp16-operand implementation may legitimately land on the other side:
  * CE   : relative gap between the reference scores ranked n_keep and n_keep+1 (any of the three
           candidate-elimination stages, any sequence) below TAU_CE;
  * argmax: relative gap between the two largest Hann-windowed responses below TAU_AM.
"""
import torch

TAU_CE = 5e-4     # ~2x the largest CE-score error measured for the fp16-operand engine (2.6e-4)
TAU_AM = 2e-3


def ce_boundary_gaps(out):
    """relative keep-boundary gaps, shape (3, B), from an oracle model_forward output."""
    gaps = []
    for sc, rem in zip(out["ce_scores"], out["removed_indexes_s"]):
        keep = sc.shape[1] - rem.shape[1]
        srt = torch.sort(sc, dim=1, descending=True).values
        gaps.append((srt[:, keep - 1] - srt[:, keep]) / srt[:, keep - 1].clamp_min(1e-30))
    return torch.stack(gaps)


def argmax_gap(response):
    """relative gap between the top-2 values of each (B, 256) response map."""
    top = response.flatten(1).topk(2, dim=1).values
    return (top[:, 0] - top[:, 1]) / top[:, 0]


def is_near_tie(out, window):
    """(B,) bool: the reference frame sits at a documented near-tie."""
    ce = (ce_boundary_gaps(out) < TAU_CE).any(dim=0)
    am = argmax_gap(window * out["score_map"]) < TAU_AM
    return ce | am


def ce_set_status(removed_gpu, removed_ref, ce_scores_ref):
    """Per sequence: 'equal' (all three kept sets identical), 'near_tie' (the first disagreeing stage
    differs only by tokens whose REFERENCE scores lie within TAU_CE of the keep threshold), or 'mismatch'.
    removed_*: three (B, n_removed) index tensors; ce_scores_ref: three (B, n_s) score tensors whose
    column order is the reference's slot order at that stage (stage 0: raster order)."""
    B = removed_ref[0].shape[0]
    out = []
    for b in range(B):
        status = "equal"
        for i in range(3):
            a = set(int(v) for v in removed_gpu[i][b].tolist())
            r = set(int(v) for v in removed_ref[i][b].tolist())
            if a == r:
                continue
            sc = ce_scores_ref[i][b]
            keep = sc.numel() - len(r)
            srt = torch.sort(sc, descending=True).values
            thr = 0.5 * (srt[keep - 1] + srt[keep]).item()
            # scores of the slots the reference kept-but-we-dropped / dropped-but-we-kept: the |a ^ r| / 2
            # closest scores on each side of the threshold
            n = len(a ^ r) // 2
            worst = max(abs(srt[keep - n].item() - thr), abs(srt[keep + n - 1].item() - thr)) / thr
            status = "near_tie" if (n <= 2 and worst < TAU_CE) else "mismatch"
            break        # later stages see different candidates once one stage differs
        out.append(status)
    return out
