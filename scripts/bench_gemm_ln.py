"""Time the LayerNorm-fold GEMM pair against the plain GEMMs (CUDA events, L2 flushed between launches).
usage: python scripts/bench_gemm_ln.py [B] [only]   (only = resid|fold|plain|prompt: a single kernel class, for ncu)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from vipt_b200 import _lib

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
only = sys.argv[2] if len(sys.argv) > 2 else ""
L = _lib.lib()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def timeit(fn, reps=10):
    ts = []
    for it in range(reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts = sorted(ts[2:])
    return ts[len(ts) // 2]


for n in (320, 244, 190, 153):
    M = B * n
    att = (torch.randn(M, 768, device="cuda") * 0.5).half()
    Wp = (torch.randn(768, 768, device="cuda") * 0.03).half()
    W1 = (torch.randn(3072, 768, device="cuda") * 0.03).half()
    W2 = (torch.randn(768, 3072, device="cuda") * 0.03).half()
    bp = torch.randn(768, device="cuda"); b1 = torch.randn(3072, device="cuda"); cs = torch.randn(3072, device="cuda")
    x = torch.randn(M, 768, device="cuda"); x16 = torch.empty(M, 768, device="cuda", dtype=torch.float16)
    d16 = torch.empty(M, 768, device="cuda", dtype=torch.float16)
    h = torch.empty(M, 3072, device="cuda", dtype=torch.float16)
    stat = torch.rand(24, M, 2, device="cuda") * 768
    sp = _lib.stream_ptr()
    res = {}
    if only in ("", "plain"):
        res["proj"] = timeit(lambda: _lib.check(L.vipt_op_gemm(_lib.ptr(att), 768, _lib.ptr(Wp), 768, M, 768, 768, 512, 0, _lib.ptr(d16), 768, _lib.ptr(bp), None, sp)))
        res["fc1"] = timeit(lambda: _lib.check(L.vipt_op_gemm(_lib.ptr(x16), 768, _lib.ptr(W1), 768, M, 3072, 768, 512, 1, _lib.ptr(h), 3072, _lib.ptr(b1), None, sp)))
        res["fc2"] = timeit(lambda: _lib.check(L.vipt_op_gemm(_lib.ptr(h), 3072, _lib.ptr(W2), 3072, M, 768, 3072, 512, 0, _lib.ptr(d16), 768, _lib.ptr(bp), None, sp)))
    if only in ("", "resid"):
        res["proj_resid_stat"] = timeit(lambda: _lib.check(L.vipt_op_gemm_resid_stat(_lib.ptr(att), 768, _lib.ptr(Wp), 768, M, 768, 768, _lib.ptr(x), _lib.ptr(x16), _lib.ptr(bp), _lib.ptr(stat), sp)))
    if only in ("", "fold"):
        res["fc1_lnfold"] = timeit(lambda: _lib.check(L.vipt_op_gemm_lnfold(_lib.ptr(x16), 768, _lib.ptr(W1), 768, M, 3072, 768, 1, _lib.ptr(h), 3072, _lib.ptr(b1), _lib.ptr(cs), _lib.ptr(stat), 24, 1e-6, None, None, sp)))
    if only in ("", "prompt"):
        colw = torch.randn(768, 20, device="cuda") * 0.1
        part = torch.empty(6, M, 20, device="cuda")
        res["fc2_resid_prompt"] = timeit(lambda: _lib.check(L.vipt_op_gemm_resid_prompt(_lib.ptr(h), 3072, _lib.ptr(W2), 3072, M, 768, 3072, _lib.ptr(x), _lib.ptr(x16), _lib.ptr(bp), _lib.ptr(colw), _lib.ptr(part), sp)))
        res["fc2_resid_stat"] = timeit(lambda: _lib.check(L.vipt_op_gemm_resid_stat(_lib.ptr(h), 3072, _lib.ptr(W2), 3072, M, 768, 3072, _lib.ptr(x), _lib.ptr(x16), _lib.ptr(bp), _lib.ptr(stat), sp)))
    print("B=%d n=%d M=%d  " % (B, n, M) + "  ".join("%s %.1f us" % kv for kv in res.items()), flush=True)
