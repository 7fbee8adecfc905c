"""Near-tie bookkeeping for the parity protocol -- TEST INFRASTRUCTURE ONLY.

A frame is a *documented near-tie* when the fp32 reference itself is within TAU of making a different
discrete decision, so an fp16-operand implementation may legitimately land on the other side:
  * CE   : relative gap between the reference scores ranked n_keep and n_keep+1 (any of the three
           candidate-elimination stages, any sequence) below TAU_CE;
  * argmax: relative gap between the two largest Hann-windowed responses below TAU_AM.
"""
import torch

TAU_CE = 1e-3
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
