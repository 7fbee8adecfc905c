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


@pytest.mark.parametrize("n_s,n_keep", [(256, 180), (180, 126), (126, 89)])
def test_ce_topk_against_reference_sort(n_s, n_keep):
    """ce_topk_kernel against attn_blocks.py:37-63 (mean over heads, descending sort, keep ceil(0.7 n)): index work,
    bit-exact.  Sequence 0 has distinct scores (the order must equal torch.sort's); sequence 1 contains groups of EXACT
    ties, across the keep boundary too: there the reference's order is unspecified (torch.sort is not stable) and the
    documented rule of this implementation is 'lower slot first' = a stable descending sort."""
    L = _lib.lib()
    dev = torch.device("cuda")
    B, H = 3, 12
    g = torch.Generator().manual_seed(n_s)
    prob = torch.rand(B, H, 256, generator=g)
    prob[1] = torch.randint(0, 6, (1, 256), generator=g).float().expand(H, 256) / 8.0       # many exact ties
    prob[2, :, n_keep - 2:n_keep + 3] = 0.5                                                   # a tie across the boundary
    gi_in = torch.stack([torch.randperm(256, generator=g) for _ in range(B)]).int()
    ce = prob.to(dev)
    gi_out = torch.full((B, 256), -7, dtype=torch.int32, device=dev)
    perm = torch.full((B, 256), -7, dtype=torch.int32, device=dev)
    removed = torch.full((B, 256), -7, dtype=torch.int32, device=dev)
    sop = torch.full((B, 256), -7, dtype=torch.int32, device=dev)
    score = torch.zeros(B, 256, device=dev)
    gi_in_d = gi_in.to(dev)
    _lib.check(L.vipt_op_ce_topk(_lib.ptr(ce), H, n_s, n_keep, _lib.ptr(gi_in_d), _lib.ptr(gi_out), _lib.ptr(perm),
                                 _lib.ptr(removed), _lib.ptr(sop), _lib.ptr(score), B, _lib.stream_ptr()), "ce_topk")
    torch.cuda.synchronize()
    ref_score = prob[:, :, :n_s].mean(dim=1)
    # the device sums the heads in order and divides by 12: same value as torch's mean up to fp32 summation order
    np.testing.assert_allclose(score.cpu()[:, :n_s].numpy(), ref_score.numpy(), rtol=2e-6)
    dscore = score.cpu()[:, :n_s]
    order = torch.sort(dscore, dim=1, descending=True, stable=True).indices          # documented tie rule
    for b in range(B):
        keep, drop = order[b, :n_keep], order[b, n_keep:]
        assert torch.equal(perm[b, :n_keep].cpu().long(), keep)
        assert torch.equal(gi_out[b, :n_keep].cpu().long(), gi_in[b].long()[keep])
        assert torch.equal(removed[b, :n_s - n_keep].cpu().long(), gi_in[b].long()[drop])
        s = sop[b].cpu().long()
        assert torch.equal(s[gi_in[b].long()[keep]], torch.arange(n_keep))
        assert (s[gi_in[b].long()[drop]] == -1).all()
    # distinct scores: exactly the reference's (unstable) sort
    ref_order = torch.sort(ref_score[0], descending=True).indices
    if torch.equal(dscore[0], ref_score[0]):
        assert torch.equal(perm[0, :n_keep].cpu().long(), ref_order[:n_keep])


@pytest.mark.parametrize("B,cin,N,G,bn", [
    (64, 768, 768, 1, 512),      # head conv1, three towers concatenated on N, 2-CTA kernel (large batch)
    (2, 768, 768, 1, 64),        # head conv1 in latency mode (single-CTA, narrow tiles)
    (5, 256, 128, 3, 128),       # conv2: three towers as groups reading their own 256 channels
    (5, 128, 64, 3, 64),         # conv3
    (3, 64, 32, 3, 32),          # conv4
])
def test_implicit_conv3x3_against_conv2d(B, cin, N, G, bn):
    """The implicit-GEMM 3x3 convolution (4-D TMA map, zero padding by out-of-bounds fill, bias + ReLU fused) against
    torch conv2d in fp32 on the same fp16-rounded operands (ViPT/lib/models/layers/head.py:8-21 with BN folded)."""
    L = _lib.lib()
    dev = torch.device("cuda")
    g = torch.Generator().manual_seed(B * 1000 + cin)
    c_total = cin * G
    feat = (torch.randn(B, 16, 16, c_total, generator=g) * 0.7).half()
    w = (torch.randn(G, N, cin, 3, 3, generator=g) / np.sqrt(9 * cin)).half()
    bias = torch.randn(G, N, generator=g) * 0.1
    wk = w.permute(0, 1, 3, 4, 2).reshape(G, N, 9 * cin).contiguous()               # K order tap * cin + c
    out = torch.full((B * 256, G * N), float("nan"), dtype=torch.float16, device=dev)
    feat_d, wk_d, bias_d = feat.to(dev), wk.to(dev), bias.to(dev)       # named: they must outlive the launch
    _lib.check(L.vipt_op_conv3x3(_lib.ptr(feat_d), c_total, cin, cin, _lib.ptr(wk_d), G, B, N, bn,
                                 _lib.ptr(out), G * N, N, _lib.ptr(bias_d), _lib.stream_ptr()), "conv3x3")
    torch.cuda.synchronize()
    x = feat.float().permute(0, 3, 1, 2)
    for gg in range(G):
        ref = torch.relu(torch.nn.functional.conv2d(x[:, gg * cin:(gg + 1) * cin], w[gg].float(), bias[gg], padding=1))
        ref = ref.permute(0, 2, 3, 1).reshape(B * 256, N)
        got = out[:, gg * N:(gg + 1) * N].float().cpu()
        err = (got - ref).abs().max().item() / ref.abs().max().item()
        assert err < 2e-3, (gg, err)
