"""Per-kernel parity through the C ABI: fused crop vs the (cv2-pinned) oracle, fused attention vs torch fp32."""
import ctypes

import numpy as np
import pytest
import torch

from oracle import crop_oracle as co
from vipt_b200 import _lib, synth

pytestmark = pytest.mark.gpu


def _expected_rows(patch_u8, S):
    """oracle crop -> (2, tokens, 768) fp16 patch-major rows."""
    t = torch.from_numpy(co.preprocess(patch_u8))[0]           # (6,S,S) fp32
    n = S // 16
    rows = t.view(2, 3, n, 16, n, 16).permute(0, 2, 4, 1, 3, 5).reshape(2, n * n, 768)
    return rows.half()


@pytest.mark.parametrize("factor,S,row_base", [(4.0, 256, 64), (2.0, 128, 0)])
def test_crop_bit_exact(factor, S, row_base):
    L = _lib.lib()
    frames, _, _ = synth.synthetic_sequence(7, 1, "rgbt")
    im = frames[0]
    H, W = im.shape[:2]
    boxes = [[250.0, 190.0, 60.0, 48.0], [-20.5, -10.25, 90.0, 70.0], [W - 40.0, H - 30.0, 80.0, 64.0],
             [300.0, 200.0, 128.0, 128.0] if S == 256 else [300.0, 200.0, 128.0, 128.0],
             [310.2, 222.7, 12.0, 10.0], [100.5, 50.5, 200.0, 300.0], [301.5, 201.5, 33.0, 31.0],
             [10.0, 400.0, 63.9, 64.1], [250.5, 190.5, 61.0, 47.0], [0.0, 0.0, float(W), float(H)]]
    B = len(boxes)
    dev = torch.device("cuda")
    fr = torch.from_numpy(im).to(dev)
    ptrs = torch.tensor([fr.data_ptr()] * B, dtype=torch.int64, device=dev)
    hw = torch.tensor([[H, W]] * B, dtype=torch.int32, device=dev)
    bx = torch.tensor(boxes, dtype=torch.float64, device=dev)
    rows_per_seq = 320
    ap = torch.zeros(2, B * rows_per_seq, 768, dtype=torch.float16, device=dev)
    geom = torch.zeros(B, 4, dtype=torch.float64, device=dev)
    err = torch.zeros(1, dtype=torch.int32, device=dev)
    _lib.check(L.vipt_op_crop(_lib.ptr(ptrs), _lib.ptr(hw), _lib.ptr(bx), factor, S, row_base, rows_per_seq,
                              _lib.ptr(ap), _lib.ptr(geom), _lib.ptr(err), B, _lib.stream_ptr()), "vipt_op_crop")
    torch.cuda.synchronize()
    assert err.item() == 0
    ntok = (S // 16) ** 2
    ap = ap.view(2, B, rows_per_seq, 768).cpu()
    for b, box in enumerate(boxes):
        patch, rf = co.sample_target(im, box, factor, S)
        x1, y1, cs = co.crop_geometry(box, factor, H, W)
        assert geom[b].tolist() == [float(x1), float(y1), float(cs), rf]
        exp = _expected_rows(patch, S)
        got = ap[:, b, row_base:row_base + ntok]
        assert torch.equal(got, exp), "box %d: %d mismatching values" % (b, (got != exp).sum().item())


def test_crop_too_small_box_sets_error():
    L = _lib.lib()
    dev = torch.device("cuda")
    fr = torch.zeros(64, 64, 6, dtype=torch.uint8, device=dev)
    ptrs = torch.tensor([fr.data_ptr()], dtype=torch.int64, device=dev)
    hw = torch.tensor([[64, 64]], dtype=torch.int32, device=dev)
    bx = torch.tensor([[1.0, 1.0, 0.0, 5.0]], dtype=torch.float64, device=dev)
    ap = torch.zeros(2, 320, 768, dtype=torch.float16, device=dev)
    geom = torch.zeros(1, 4, dtype=torch.float64, device=dev)
    err = torch.zeros(1, dtype=torch.int32, device=dev)
    L.vipt_op_crop(_lib.ptr(ptrs), _lib.ptr(hw), _lib.ptr(bx), 4.0, 256, 64, 320, _lib.ptr(ap), _lib.ptr(geom),
                   _lib.ptr(err), 1, _lib.stream_ptr())
    torch.cuda.synchronize()
    assert err.item() == -5          # VIPT_ERR_SMALL_BOX <-> Exception('Too small bounding box.')


@pytest.mark.parametrize("B,N", [(2, 320), (3, 244), (2, 190), (5, 153), (1, 64),
                                 (13, 320), (14, 244), (13, 190), (15, 153),    # B * 12 >= #SMs: persistent kernel,
                                 (64, 320), (64, 153)])                         # work split inside (sequence, head) items
def test_attention(B, N):
    L = _lib.lib()
    g = torch.Generator(device="cuda").manual_seed(N)
    qkv = (torch.randn(B * N, 2304, device="cuda", generator=g) * 1.5).half()
    out = torch.full((B * N, 768), float("nan"), device="cuda", dtype=torch.float16)
    ce = torch.full((B, 12, 256), float("nan"), device="cuda", dtype=torch.float32)
    _lib.check(L.vipt_op_attention(_lib.ptr(qkv), B, N, _lib.ptr(out), _lib.ptr(ce), _lib.stream_ptr()),
               "vipt_op_attention")
    torch.cuda.synchronize()
    q, k, v = qkv.float().view(B, N, 3, 12, 64).permute(2, 0, 3, 1, 4)
    p = torch.softmax((q @ k.transpose(-1, -2)) * 0.125, dim=-1)
    ref = (p @ v).transpose(1, 2).reshape(B * N, 768)
    err = (out.float() - ref).abs().max().item()
    assert err < 4e-3 * max(1.0, ref.abs().max().item()), err
    if N > 64:
        ce_ref = p[:, :, 27, 64:]
        ce_err = (ce[:, :, :N - 64] - ce_ref).abs().max().item()
        assert ce_err < 1e-5 + 1e-4 * ce_ref.max().item(), ce_err
