"""init + N un-graphed tracking steps at batch B (for ncu captures: 111 launches per step, 1 at init)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from vipt_b200 import synth
from vipt_b200.tracker import BatchedViPTTracker
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
N = int(sys.argv[2]) if len(sys.argv) > 2 else 3
graph = len(sys.argv) > 3 and sys.argv[3] == "graph"
sd = synth.synthetic_state_dict("vipt_deep", 0)
base = [synth.synthetic_sequence(1000 + s, N + 1, "rgbt") for s in range(min(B, 8))]
trk = BatchedViPTTracker(sd, B, "vipt_deep", use_graph=graph)
dev = trk.engine.device
frames = [[torch.from_numpy(base[i % len(base)][0][t]).to(dev) for i in range(B)] for t in range(N + 1)]
trk.initialize(frames[0], [base[i % len(base)][1] for i in range(B)])
for t in range(1, N + 1):
    boxes, scores = trk.track(frames[t])
print("ok", boxes[0], scores[0])
