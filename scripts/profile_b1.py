"""per-kernel-class device time of one un-graphed step at batch B (default 1)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from vipt_b200 import synth
from vipt_b200.tracker import BatchedViPTTracker
B = int(sys.argv[1]) if len(sys.argv) > 1 else 1
sd = synth.synthetic_state_dict("vipt_deep", 0)
fr, box0, _ = synth.synthetic_sequence(1000, 4, "rgbt")
trk = BatchedViPTTracker(sd, B, "vipt_deep", use_graph=False)
dev = trk.engine.device
f = [[torch.from_numpy(fr[t]).to(dev) for _ in range(B)] for t in range(4)]
trk.initialize(f[0], [box0] * B)
ptrs = [torch.tensor([x.data_ptr() for x in f[t]], dtype=torch.int64).to(dev) for t in range(4)]
for t in (1, 2, 3):
    p = trk.engine.profile_step(ptrs[t])
print({k: (round(v[0] * 1e3, 1), v[1]) for k, v in p.items()}, "sum us", round(sum(v[0] for v in p.values()) * 1e3, 1))
