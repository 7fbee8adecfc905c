"""Compare call patterns of the host-frame API at batch B: synchronous track(), pipelined track_async()+collect(),
and the host-side cost of one enqueue.  usage: python scripts/e2e_modes.py [B] [steps]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from vipt_b200 import synth
from vipt_b200.tracker import BatchedViPTTracker
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
K = int(sys.argv[2]) if len(sys.argv) > 2 else 60
RING = 8
sd = synth.synthetic_state_dict("vipt_deep", 0)
base = [synth.synthetic_sequence(1000 + s, RING + 1, "rgbt") for s in range(8)]
host = [[torch.from_numpy(np.ascontiguousarray(base[i % 8][0][t])).pin_memory() for i in range(B)] for t in range(RING + 1)]
trk = BatchedViPTTracker(sd, B, "vipt_deep", use_graph=True)
trk.initialize(host[0], [base[i % 8][1] for i in range(B)])
for k in range(5):
    trk.track(host[1 + k % RING])
def sync_mode():
    t0 = time.perf_counter()
    for k in range(K):
        trk.track(host[1 + k % RING])
    return (time.perf_counter() - t0) / K * 1e3
def pipe_mode():
    t0 = time.perf_counter()
    trk.track_async(host[1])
    for k in range(1, K):
        trk.track_async(host[1 + k % RING]); trk.collect()
    trk.collect()
    return (time.perf_counter() - t0) / K * 1e3
def host_cost():
    torch.cuda.synchronize()
    ts = []
    for k in range(3):
        t0 = time.perf_counter(); trk.track_async(host[1 + k]); ts.append((time.perf_counter() - t0) * 1e3)
    for k in range(3):
        t0 = time.perf_counter(); trk.collect(); ts.append((time.perf_counter() - t0) * 1e3)
    return ts
for rep in range(2):
    print("sync %.3f ms/step   pipelined %.3f ms/step" % (sync_mode(), pipe_mode()))
print("host cost of 3 enqueues then 3 collects (ms):", ["%.3f" % t for t in host_cost()])
