"""BASELINE.json configs[4]: throughput / latency sweep over the number of concurrent sequences per B200
(B = 1 .. 128; 128 per GPU = the 1024-sequence / 8-GPU point), ViPT-deep RGBT, frames resident in HBM, CUDA-graph
replay, CUDA events on the engine's stream.  Ring of 8 frame sets; beyond B = 32 that is larger than L2.
usage: python scripts/batch_sweep.py [out.json] [steps]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from vipt_b200 import synth
from vipt_b200.tracker import BatchedViPTTracker
out = sys.argv[1] if len(sys.argv) > 1 else None
K = int(sys.argv[2]) if len(sys.argv) > 2 else 40
RING, ND = 8, 8
sd = synth.synthetic_state_dict("vipt_deep", 0)
base = [synth.synthetic_sequence(1000 + s, RING + 1, "rgbt") for s in range(ND)]
rows = []
ref_box = None
BS = [int(b) for b in sys.argv[3].split(",")] if len(sys.argv) > 3 else [1, 2, 4, 8, 16, 32, 64, 96, 128]
for B in BS:
    trk = BatchedViPTTracker(sd, B, "vipt_deep", use_graph=True)
    eng = trk.engine
    dev = eng.device
    with torch.cuda.stream(eng.stream):
        devf = [[torch.from_numpy(base[i % ND][0][t]).to(dev) for i in range(B)] for t in range(RING + 1)]
        ptrs = [torch.tensor([f.data_ptr() for f in devf[t]], dtype=torch.int64).to(dev) for t in range(RING + 1)]
    eng.stream.synchronize()
    trk.initialize(devf[0], [base[i % ND][1] for i in range(B)])
    eng.track_enqueue(ptrs[1], True)
    b1 = eng.results()[0][0].copy()          # slot 0 after the first tracked frame: must not depend on B
    if ref_box is None:
        ref_box = b1
    for k in range(8):
        eng.track_enqueue(ptrs[1 + (k + 1) % RING], True)
    eng.results()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(K + 1)]
    with torch.cuda.stream(eng.stream):
        ev[0].record()
        for k in range(K):
            eng.track_enqueue(ptrs[1 + (k + 9) % RING], True)
            ev[k + 1].record()
    eng.stream.synchronize()
    per = sorted(ev[k].elapsed_time(ev[k + 1]) for k in range(K))
    total = ev[0].elapsed_time(ev[K])
    row = {"sequences": B, "ms_per_step": total / K, "p50_ms": per[K // 2], "p90_ms": per[(9 * K) // 10],
           "frames_per_s": B * K / total * 1e3, "slot0_box_dev_px": float(np.abs(b1 - ref_box).max())}
    rows.append(row)
    print(json.dumps(row), flush=True)
    del trk, eng, devf, ptrs
    torch.cuda.empty_cache()
if out:
    json.dump({"workload": "ViPT-deep RGBT 630x480x6, frames in HBM, graph replay, %d timed steps per point" % K, "points": rows},
              open(out, "w"), indent=1)
