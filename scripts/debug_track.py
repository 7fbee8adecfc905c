import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import vipt_oracle as vo, crop_oracle as co
from vipt_b200 import synth
from vipt_b200.tracker import BatchedViPTTracker
sd = synth.synthetic_state_dict("vipt_deep", 0)
frames, box0, _ = synth.synthetic_sequence(3, 3, "rgbd")
ref = vo.OracleTracker(sd, "vipt_deep"); ref.initialize(frames[0], {"init_bbox": list(box0)})
trk = BatchedViPTTracker(sd, 1, "vipt_deep", use_graph=False); trk.initialize([frames[0]], [box0])
eng = trk.engine
def rows(patch, S):
    t = torch.from_numpy(co.preprocess(patch))[0]; n = S // 16
    return t.view(2, 3, n, 16, n, 16).permute(0, 2, 4, 1, 3, 5).reshape(2, n * n, 768).half()
ap = eng.buffer("apatch").cpu()
print("template rows equal:", torch.equal(ap[:, 0, :64], rows(ref.z_patch, 128)))
r = ref.track(frames[1]); boxes, scores = trk.track([frames[1]])
ap = eng.buffer("apatch").cpu()
print("search rows equal:", torch.equal(ap[:, 0, 64:], rows(ref.last["x_patch"], 256)), "template still:", torch.equal(ap[:, 0, :64], rows(ref.z_patch, 128)))
out = ref.last["out"]
for k in ("score_map", "size_map", "offset_map"):
    a = eng.buffer(k).cpu(); print(k, "rel", ((a - out[k]).abs().max() / out[k].abs().max()).item())
resp_ref = (ref.window * out["score_map"]).flatten(); resp = (ref.window * eng.buffer("score_map").cpu()).flatten()
print("argmax ref", resp_ref.argmax().item(), "gpu", resp.argmax().item(), "top2 ref", resp_ref.topk(2))
print("geom", eng.buffer("geom_x").cpu().tolist(), "rf ref", ref.last["resize_factor"])
print("gpu", boxes, scores, "ref", r)
rem = eng.buffer("removed").cpu(); n = [76, 54, 37]
print([set(rem[i, 0, :n[i]].tolist()) == set(out["removed_indexes_s"][i][0].tolist()) for i in range(3)])
