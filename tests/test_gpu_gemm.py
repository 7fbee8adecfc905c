"""tcgen05 GEMM through the C ABI vs a torch fp32 reference on the same fp16-rounded operands."""
import pytest
import torch

from vipt_b200 import _lib

pytestmark = pytest.mark.gpu


def _ref(A, W, bias, epi, resid):
    y = A.float() @ W.float().t() + bias
    if epi == 1:
        y = torch.nn.functional.gelu(y)
    if epi == 2:
        y = torch.relu(y)
    if epi == 4:
        y = y + resid
    return y


@pytest.mark.parametrize("M,N,K,bn,epi", [
    (128, 256, 64, 256, 3), (128, 256, 128, 256, 3), (256, 256, 768, 256, 3),
    (320, 768, 768, 256, 0), (1000, 2304, 768, 256, 0), (777, 3072, 768, 256, 1), (640, 768, 3072, 256, 4),
    (256, 256, 64, 512, 0), (256, 256, 768, 512, 0), (1000, 2304, 768, 512, 0), (777, 3072, 768, 512, 1),
    (20480, 768, 3072, 512, 0), (9792, 768, 768, 512, 2), (130, 256, 128, 512, 1),
    (20480, 768, 3072, 640, 0), (15616, 768, 768, 640, 0), (300, 128, 64, 640, 1), (777, 384, 768, 640, 2),   # 128-wide pair tiles
    (1000, 768, 768, 512, 3), (20480, 768, 768, 512, 3), (300, 128, 64, 640, 3),                              # 2-CTA kernel, fp32 out
    (15616, 768, 768, 512, 0), (5000, 3072, 768, 512, 1), (19000, 768, 3072, 512, 0),      # last round cut into column strips
    (500, 128, 2304, 128, 2), (300, 64, 1152, 64, 2), (260, 32, 576, 32, 2), (20480, 768, 768, 256, 4),
])
def test_gemm(M, N, K, bn, epi):
    L = _lib.lib()
    g = torch.Generator(device="cuda").manual_seed(M + N + K)
    A = (torch.randn(M, K, device="cuda", generator=g) * 0.5).half()
    W = (torch.randn(N, K, device="cuda", generator=g) * 0.05).half()
    bias = torch.randn(N, device="cuda", generator=g)
    resid = torch.randn(M, N, device="cuda", generator=g) if epi == 4 else None
    out = torch.full((M, N), float("nan"), device="cuda", dtype=torch.float32 if epi in (3, 4) else torch.float16)
    _lib.check(L.vipt_op_gemm(_lib.ptr(A), K, _lib.ptr(W), K, M, N, K, bn, epi, _lib.ptr(out), N, _lib.ptr(bias),
                              _lib.ptr(resid), _lib.stream_ptr()), "vipt_op_gemm")
    torch.cuda.synchronize()
    ref = _ref(A, W, bias, epi, resid)
    err = (out.float() - ref).abs().max().item()
    tol = 2e-3 * max(1.0, ref.abs().max().item()) if out.dtype == torch.float16 else 1e-3
    if not err < tol:
        bad = ((out.float() - ref).abs() > tol).nonzero()
        print("bad count", bad.shape[0], "first", bad[:8].tolist(), "last", bad[-8:].tolist())
        print("out", out[:2, :8], "ref", ref[:2, :8])
    assert err < tol, (err, tol)


@pytest.mark.parametrize("M,N,K", [(256, 256, 64), (300, 768, 768), (20480, 768, 768), (9792, 768, 3072), (130, 512, 128),
                                   (15616, 768, 768), (19000, 768, 3072)])      # the last two: tail in column strips
def test_gemm_resid_stat(M, N, K):
    """LayerNorm fold, producer (gemm2.cu EPI_RESID_STAT): x += A W^T + b in place (fp32), fp16 copy, per-row partial sums."""
    L = _lib.lib()
    g = torch.Generator(device="cuda").manual_seed(M + N + K)
    A = (torch.randn(M, K, device="cuda", generator=g) * 0.5).half()
    W = (torch.randn(N, K, device="cuda", generator=g) * 0.05).half()
    bias = torch.randn(N, device="cuda", generator=g)
    x0 = torch.randn(M, N, device="cuda", generator=g) * 2 + 0.3
    x = x0.clone()
    x16 = torch.full((M, N), float("nan"), device="cuda", dtype=torch.float16)
    parts = N // 32
    stat = torch.full((parts, M, 2), float("nan"), device="cuda")
    _lib.check(L.vipt_op_gemm_resid_stat(_lib.ptr(A), K, _lib.ptr(W), K, M, N, K, _lib.ptr(x), _lib.ptr(x16), _lib.ptr(bias),
                                         _lib.ptr(stat), _lib.stream_ptr()), "vipt_op_gemm_resid_stat")
    torch.cuda.synchronize()
    ref = x0 + A.float() @ W.float().t() + bias
    assert (x - ref).abs().max().item() < 1e-3
    assert torch.equal(x16, x.half())                                # the fp16 copy is the rounding of the stored fp32 value
    # part 2 g + h = the 16-column groups h and h + 2 of columns [64 g, 64 g + 64) (one epilogue thread's columns)
    xp = x.view(M, N // 64, 2, 2, 16).permute(0, 1, 3, 2, 4).reshape(M, parts, 32)
    assert (stat[..., 0].t() - xp.sum(-1)).abs().max().item() < 2e-3
    assert ((stat[..., 1].t() - (xp * xp).sum(-1)).abs() / (xp * xp).sum(-1)).max().item() < 1e-5


@pytest.mark.parametrize("M,N,gelu", [(256, 256, 0), (300, 2304, 0), (20480, 3072, 1), (9792, 3072, 1), (130, 512, 1),
                                         (15616, 2304, 0)])        # last round in column strips
def test_gemm_lnfold(M, N, gelu):
    """LayerNorm fold, consumer (EPI_LNFOLD_*): with W' = fp16(gamma W0), colsum(W'), bias' = b0 + W0 beta and the row
    statistics of x, the GEMM over fp16(x) is LayerNorm(x) W0^T + b0 (attn_blocks.py:94,103) to fp16-operand accuracy."""
    L = _lib.lib()
    K = 768
    g = torch.Generator(device="cuda").manual_seed(M + N)
    x = torch.randn(M, K, device="cuda", generator=g) * (0.5 + 3 * torch.rand(M, 1, device="cuda", generator=g)) + \
        0.3 * torch.randn(M, 1, device="cuda", generator=g)
    gamma = 1 + 0.2 * torch.randn(K, device="cuda", generator=g)
    beta = 0.1 * torch.randn(K, device="cuda", generator=g)
    W0 = torch.randn(N, K, device="cuda", generator=g) * 0.03
    b0 = torch.randn(N, device="cuda", generator=g)
    Wf = (W0 * gamma).half()
    colsum = Wf.float().sum(1).contiguous()
    bias = (b0 + W0 @ beta).contiguous()
    x16 = x.half()
    xp = x.view(M, 24, 32)
    stat = torch.stack([xp.sum(-1), (xp * xp).sum(-1)], -1).permute(1, 0, 2).contiguous()      # (24, M, 2)
    out = torch.full((M, N), float("nan"), device="cuda", dtype=torch.float16)
    _lib.check(L.vipt_op_gemm_lnfold(_lib.ptr(x16), K, _lib.ptr(Wf), K, M, N, K, gelu, _lib.ptr(out), N, _lib.ptr(bias),
                                     _lib.ptr(colsum), _lib.ptr(stat), 24, 1e-6, None, None, _lib.stream_ptr()),
               "vipt_op_gemm_lnfold")
    torch.cuda.synchronize()
    ref = torch.nn.functional.layer_norm(x.double(), (K,), gamma.double(), beta.double(), 1e-6) @ W0.double().t() + b0.double()
    if gelu:
        ref = torch.nn.functional.gelu(ref)
    err = (out.double() - ref).abs().max().item()
    assert err < 4e-3 * max(1.0, ref.abs().max().item()), err


@pytest.mark.parametrize("M,N", [(300, 2304), (20480, 2304), (15616, 768)])
def test_gemm_lnfold_final_stats_and_extra_kblock(M, N):
    """Prompt fold, consumer: final row statistics (stat_parts = 0) and the rank-8 update as one extra k-block:
    out = rstd (x16 W'^T + f (W'U)^T) - mean rstd colsum + bias, which is LayerNorm(x16 + U f) W0^T + b0 when the
    statistics are those of x16 + U f (vit_ce_prompt.py:308-310 followed by attn_blocks.py:94)."""
    L = _lib.lib()
    K = 768
    g = torch.Generator(device="cuda").manual_seed(M + N + 1)
    v = torch.randn(M, K, device="cuda", generator=g) * 1.5
    U = torch.randn(K, 8, device="cuda", generator=g) * 0.05
    f = torch.randn(M, 8, device="cuda", generator=g) * 2
    gamma = 1 + 0.2 * torch.randn(K, device="cuda", generator=g)
    beta = 0.1 * torch.randn(K, device="cuda", generator=g)
    W0 = torch.randn(N, K, device="cuda", generator=g) * 0.03
    b0 = torch.randn(N, device="cuda", generator=g)
    Wf = (W0 * gamma).half()
    colsum = Wf.float().sum(1).contiguous()
    bias = (b0 + W0 @ beta).contiguous()
    x = v.half().float() + f.half().float() @ U.t()                       # what the fp16 operands represent
    mean, var = x.mean(1), x.var(1, unbiased=False)
    rs = (var + 1e-6).rsqrt()
    stat = torch.stack([rs, -mean * rs], -1).contiguous()
    Ax = torch.zeros(M, 64, device="cuda", dtype=torch.float16); Ax[:, :8] = f.half()
    Wx = torch.zeros(N, 64, device="cuda", dtype=torch.float16); Wx[:, :8] = (Wf.float() @ U).half()
    out = torch.full((M, N), float("nan"), device="cuda", dtype=torch.float16)
    _lib.check(L.vipt_op_gemm_lnfold(_lib.ptr(v.half()), K, _lib.ptr(Wf), K, M, N, K, 0, _lib.ptr(out), N, _lib.ptr(bias),
                                     _lib.ptr(colsum), _lib.ptr(stat), 0, 1e-6, _lib.ptr(Ax), _lib.ptr(Wx), _lib.stream_ptr()),
               "vipt_op_gemm_lnfold")
    torch.cuda.synchronize()
    ref = torch.nn.functional.layer_norm(x.double(), (K,), gamma.double(), beta.double(), 1e-6) @ W0.double().t() + b0.double()
    err = (out.double() - ref).abs().max().item()
    assert err < 4e-3 * max(1.0, ref.abs().max().item()), err


@pytest.mark.parametrize("M,K", [(300, 3072), (20480, 3072), (9792, 3072), (1000, 768)])
def test_gemm_resid_prompt(M, K):
    """Prompt fold, producer (EPI_RESID_PROMPT): residual add in place, fp16 copy and the 19 row sums per (tile, half)."""
    L = _lib.lib()
    N = 768
    g = torch.Generator(device="cuda").manual_seed(M + K)
    A = (torch.randn(M, K, device="cuda", generator=g) * 0.5).half()
    W = (torch.randn(N, K, device="cuda", generator=g) * 0.03).half()
    bias = torch.randn(N, device="cuda", generator=g)
    colw = torch.zeros(N, 20, device="cuda")
    colw[:, :17] = torch.randn(N, 17, device="cuda", generator=g) * 0.1
    x0 = torch.randn(M, N, device="cuda", generator=g) * 2 + 0.3
    x = x0.clone()
    x16 = torch.full((M, N), float("nan"), device="cuda", dtype=torch.float16)
    part = torch.full((6, M, 20), float("nan"), device="cuda")
    _lib.check(L.vipt_op_gemm_resid_prompt(_lib.ptr(A), K, _lib.ptr(W), K, M, N, K, _lib.ptr(x), _lib.ptr(x16), _lib.ptr(bias),
                                           _lib.ptr(colw), _lib.ptr(part), _lib.stream_ptr()), "vipt_op_gemm_resid_prompt")
    torch.cuda.synchronize()
    ref = x0 + A.float() @ W.float().t() + bias
    assert (x - ref).abs().max().item() < 2e-3
    assert torch.equal(x16, x.half())
    # part 2 t + h = the 16-column groups h, h + 2, .. of tile t; sums over those 128 columns
    xp = x.double().view(M, 3, 8, 2, 16).permute(1, 3, 0, 2, 4).reshape(6, M, 128)
    cw = colw.double().view(3, 8, 2, 16, 20).permute(0, 2, 1, 3, 4).reshape(6, 128, 20)
    want = torch.zeros(6, M, 20, dtype=torch.float64, device="cuda")
    want[..., 0] = xp.sum(-1)
    want[..., 1] = (xp * xp).sum(-1)
    want[..., 2:19] = torch.einsum("pmc,pck->pmk", xp, cw[..., :17])
    scale = want.abs().amax(dim=(0, 1)).clamp_min(1.0)
    assert ((part.double() - want).abs().amax(dim=(0, 1)) / scale).max().item() < 1e-4
