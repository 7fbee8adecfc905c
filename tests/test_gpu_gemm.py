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
