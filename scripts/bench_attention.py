"""Time vipt_op_attention alone (CUDA events, L2 flushed between launches).  usage: python scripts/bench_attention.py [B]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from vipt_b200 import _lib

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
L = _lib.lib()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for N in (320, 244, 190, 153):
    qkv = (torch.randn(B * N, 2304, device="cuda") * 1.5).half()
    out = torch.empty(B * N, 768, device="cuda", dtype=torch.float16)
    ce = torch.empty(B, 12, 256, device="cuda", dtype=torch.float32)
    ts = []
    for it in range(12):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _lib.check(L.vipt_op_attention(_lib.ptr(qkv), B, N, _lib.ptr(out), _lib.ptr(ce), _lib.stream_ptr()), "att")
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts = sorted(ts[2:])
    print("diag=%s B=%d N=%d  median %.1f us  min %.1f us" % (os.environ.get("VIPT_AT2_DIAG", "0"), B, N, ts[len(ts) // 2], ts[0]))
