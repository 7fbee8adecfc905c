"""Compare intermediate buffers of the engine with and without the LayerNorm / prompt folds (large-batch paths).
usage: python scripts/debug_fold.py [B]   -- stages: 3 + 7 i: x / xn entering block i, 4 + 7 i: qkv, 6 + 7 i: after proj,
8 + 7 i: hbuf, 9 + 7 i: after fc2"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from vipt_b200 import synth
from vipt_b200.engine import NativeEngine

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
sd = synth.weight_set("peaked", "vipt_deep")
g = torch.Generator().manual_seed(5)
z = torch.randn(B, 6, 128, 128, generator=g).cuda()
x = torch.randn(B, 6, 256, 256, generator=g).cuda()
os.environ["VIPT_NO_PROMPT_FOLD"] = "1"
os.environ["VIPT_NO_LNFOLD"] = "1"
ref = NativeEngine(sd, B, "vipt_deep")
del os.environ["VIPT_NO_LNFOLD"]
if os.environ.get("KEEP_NO_PROMPT_FOLD") is None:
    del os.environ["VIPT_NO_PROMPT_FOLD"]
new = NativeEngine(sd, B, "vipt_deep")
print("lnfold layers", new.info("lnfold_layers"), "promptfold layers", new.info("promptfold_layers"))


def rel(a, b):
    return ((a.float() - b.float()).abs().max() / b.float().abs().max().clamp_min(1e-9)).item()


WIDTH = {"qkv": 2304, "hbuf": 3072, "xA": 768, "xn": 768}
for i in range(1, 3):      # blocks before the first candidate elimination: the residual stream is xA, 320 tokens
    # comparable states: qkv (stage 4 + 7 i), x after attn.proj and norm2 (7 + 7 i), fc1 output (8 + 7 i)
    for st, names in ((4 + 7 * i, ("qkv",)), (7 + 7 * i, ("xA",)), (8 + 7 * i, ("hbuf",))):
        ref.forward(z, x, debug_stop=st)
        new.forward(z, x, debug_stop=st)
        out = []
        for nm in names:
            n = B * 320 * WIDTH[nm]
            a, b_ = new.buffer(nm).flatten()[:n].view(B * 320, -1), ref.buffer(nm).flatten()[:n].view(B * 320, -1)
            d = (a.float() - b_.float()).abs()
            out.append("%s rel %.2e (worst row %d col %d)" % (nm, rel(a, b_), int(d.max(1).values.argmax()), int(d.max(0).values.argmax())))
        print("layer %d stage %d: %s" % (i, st, "  ".join(out)), flush=True)
ref.forward(z, x)
new.forward(z, x)
print("final score_map rel", rel(new.buffer("score_map"), ref.buffer("score_map")), "removed equal",
      bool((new.buffer("removed") == ref.buffer("removed")).all()))
