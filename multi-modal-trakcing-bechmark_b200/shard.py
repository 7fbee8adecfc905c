"""Sequence sharding across the GPUs of one box.

The reference fans sequences out over GPUs with a process pool and no communication
(ViPT/RGBT_workspace/test_rgbt_mgpus.py:80-84,180-184: `gpu = worker_id % num_gpu`).  Here: one process
per GPU, sequences assigned greedily by frame count (longest first, to the least-loaded rank), no
collective on the hot path, and one gather of the per-sequence result boxes at the end.
"""
import numpy as np
import torch
import torch.distributed as dist


def assign_sequences(frame_counts, world):
    """-> list (len world) of lists of sequence indices; deterministic, balanced by total frames."""
    order = sorted(range(len(frame_counts)), key=lambda i: (-int(frame_counts[i]), i))
    load = [0] * world
    out = [[] for _ in range(world)]
    for i in order:
        r = min(range(world), key=lambda k: (load[k], k))
        out[r].append(i)
        load[r] += int(frame_counts[i])
    return [sorted(v) for v in out]


def gather_boxes(local_ids, local_boxes, n_sequences, device=None):
    """Final gather: every rank contributes {sequence id -> (T, 4) boxes}; rank 0 (and everyone) gets
    the full list ordered by sequence id.  Boxes are padded to the longest sequence for the collective
    (a few KB per sequence; NVLink bandwidth is irrelevant here)."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return {i: np.asarray(b, np.float64) for i, b in zip(local_ids, local_boxes)}
    world = dist.get_world_size()
    dev = device if device is not None else torch.device("cpu")
    t_max = torch.tensor([max([len(b) for b in local_boxes] + [0])], dtype=torch.int64, device=dev)
    n_loc = torch.tensor([len(local_ids)], dtype=torch.int64, device=dev)
    dist.all_reduce(t_max, op=dist.ReduceOp.MAX)
    n_max = n_loc.clone()
    dist.all_reduce(n_max, op=dist.ReduceOp.MAX)
    T, N = int(t_max.item()), int(n_max.item())
    payload = torch.full((N, T + 1, 4), float("nan"), dtype=torch.float64, device=dev)
    for k, (i, b) in enumerate(zip(local_ids, local_boxes)):
        b = torch.as_tensor(np.asarray(b, np.float64), device=dev).reshape(-1, 4)
        payload[k, 0, 0] = float(i); payload[k, 0, 1] = float(b.shape[0])
        payload[k, 1:1 + b.shape[0]] = b
    bufs = [torch.empty_like(payload) for _ in range(world)]
    dist.all_gather(bufs, payload)
    out = {}
    for buf in bufs:
        buf = buf.cpu().numpy()
        for k in range(N):
            if np.isnan(buf[k, 0, 0]):
                continue
            i, t = int(buf[k, 0, 0]), int(buf[k, 0, 1])
            out[i] = buf[k, 1:1 + t].copy()
    assert len(out) == n_sequences, "gathered %d of %d sequences" % (len(out), n_sequences)
    return out
