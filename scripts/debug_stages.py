"""Stage-by-stage comparison of the native engine against the CPU oracle (debug aid, run on a B200)."""
import sys
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import vipt_oracle as vo
from vipt_b200 import synth
from vipt_b200.engine import NativeEngine

ptype = sys.argv[1] if len(sys.argv) > 1 else "vipt_deep"
B = 2
sd = synth.synthetic_state_dict(ptype, 0)
g = torch.Generator().manual_seed(123)
z = torch.randn(B, 6, 128, 128, generator=g); x = torch.randn(B, 6, 256, 256, generator=g)
mask = vo.ce_template_mask(B)
trace = {}
with torch.no_grad():
    ref = vo.model_forward(sd, z, x, mask, ptype, trace=trace)
eng = NativeEngine(sd, B, ptype)


def rel(a, b):
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-9)).item()


ntok = [eng.info("n_tok_%d" % i) for i in range(13)]
print("n_tok", ntok)
for i in range(12):
    eng.forward(z, x, debug_stop=3 + 7 * i)
    torch.cuda.synchronize()
    n = ntok[i]
    buf = "xA"
    xa = eng.buffer("xA").view(-1)[:B * n * 768].view(B, n, 768).cpu()
    xb = eng.buffer("xB").view(-1)[:B * n * 768].view(B, n, 768).cpu()
    r = trace["x_in_%d" % i]
    print("layer %2d x_in: rel err xA %.2e xB %.2e  (ref absmax %.3f)" % (i, rel(xa, r), rel(xb, r), r.abs().max().item()))
eng.forward(z, x, want_feat=True)
torch.cuda.synchronize()
print("feat32 rel", rel(eng.buffer("feat32").cpu(), ref["backbone_feat"]))
for k in ("score_map", "size_map", "offset_map"):
    print(k, "rel", rel(eng.buffer(k).cpu(), ref[k]))
print("pred_boxes", eng.buffer("pred_boxes").cpu().tolist(), ref["pred_boxes"].view(-1, 4).tolist())
nrem = [76, 54, 37]
rem = eng.buffer("removed").cpu()
for i in range(3):
    a = rem[i, :, :nrem[i]].long(); b = ref["removed_indexes_s"][i]
    print("CE %d: set equal %s, order equal %s" % (i, [set(a[j].tolist()) == set(b[j].tolist()) for j in range(B)], torch.equal(a, b)))
    from oracle import near_tie
    print("   ref boundary gaps", near_tie.ce_boundary_gaps(ref)[i].tolist(), "diff tokens", [sorted(set(a[j].tolist()) ^ set(b[j].tolist())) for j in range(B)])
    sc = eng.buffer("ce_score").cpu()[i]
    rs = ref["ce_scores"][i]
    print("   ce score rel err", rel(sc[:, :rs.shape[1]], rs))
