"""Model-level map errors against the CPU oracle with and without the LayerNorm / prompt folds on the same samples.
usage: python scripts/fold_error_ab.py <weights: peaked|flat|outlier> <batch>"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
from oracle import parity, vipt_oracle as vo
from vipt_b200 import synth
import test_gpu_model as T
weights = sys.argv[1]; B = int(sys.argv[2])
gen = torch.Generator().manual_seed(23)
z = torch.randn(B, 6, 128, 128, generator=gen); x = torch.randn(B, 6, 256, 256, generator=gen)
sd = synth.weight_set(weights, "vipt_deep")
mask = vo.ce_template_mask(B)
with torch.no_grad():
    ref = vo.model_forward(sd, z, x, mask, "vipt_deep")
for mode in ("fold", "nofold"):
    if mode == "nofold": os.environ["VIPT_NO_LNFOLD"] = "1"
    m = T._model("vipt_deep", weights)
    out = m.forward(z.cuda(), x.cuda(), mask.cuda())
    eng = m.engine(B)
    removed = np.zeros((3, B, 256), np.int64)
    for s in range(3):
        r = out["removed_indexes_s"][s].cpu().numpy(); removed[s, :, :r.shape[1]] = r
    ce_score = eng.buffer("ce_score").cpu().numpy()
    gidx = parity.gpu_stage_index(ce_score, removed)
    status = [parity._ce_compare(ref, removed, ce_score, gidx, b, parity.TAU_CE_BY_SET[weights]) for b in range(B)]
    redo = [b for b in range(B) if status[b][0] == "near_tie"]
    forced = None
    if redo:
        fr = [torch.from_numpy(removed[s, redo, :parity.N_S[s] - parity.KEEP[s]]) for s in range(3)]
        with torch.no_grad():
            forced = vo.model_forward(sd, z[redo], x[redo], mask[redo], "vipt_deep", forced_removed=fr)
    errs = []
    for b in range(B):
        if status[b][0] == "mismatch": errs.append(("mismatch", b)); continue
        r, k = (ref, b) if status[b][0] == "equal" else (forced, redo.index(b))
        e = max((out[n].cpu()[b] - r[n][k]).abs().max().item() / r[n][k].abs().max().item() for n in ("score_map", "size_map", "offset_map"))
        errs.append(e)
    nums = [e for e in errs if not isinstance(e, tuple)]
    print(mode, "folds", eng.info("lnfold_layers"), eng.info("promptfold_layers"), "near_tie", len(redo), "mismatch", len(errs) - len(nums),
          "map err max %.4f p50 %.4f" % (max(nums), sorted(nums)[len(nums) // 2]), "top3", [round(v, 4) for v in sorted(nums)[-3:]], flush=True)
