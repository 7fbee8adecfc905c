"""Batched benchmark drivers: the reference's per-sequence loops
(ViPT/RGBT_workspace/test_rgbt_mgpus.py:20-117, ViPT/RGBE_workspace/test_rgbe_mgpus.py:18-84) run as ONE
lock-step loop over many sequences per GPU, reading the same dataset directory layouts and writing
byte-compatible result files, so the external LasHeR / RGBT234 / VisEvent toolkits can score them.

Differences from the reference, all on the host side: one engine per GPU instead of one model per
sequence, B sequences advance together (a finished sequence idles on its last frame), and sequences are
sharded over ranks with `shard.assign_sequences` instead of a process pool.
"""
import os

import numpy as np

from . import shard


def gen_config_rgbt(seq_path, set_type):
    """-> (rgb paths, x paths, gt (T,4) xywh).  Layouts of test_rgbt_mgpus.py:20-63."""
    def imgs(sub, ext):
        d = os.path.join(seq_path, sub)
        return sorted([seq_path + '/' + sub + '/' + p for p in os.listdir(d) if os.path.splitext(p)[1] == ext])

    if set_type == 'RGBT234' or set_type in ('LasHeR', 'lasher'):
        rgb, x = imgs('visible', '.jpg'), imgs('infrared', '.jpg')
        gt = np.loadtxt(seq_path + '/visible.txt', delimiter=',')
    elif set_type == 'GTOT':
        rgb, x = imgs('v', '.png'), imgs('i', '.png')
        c = np.loadtxt(seq_path + '/groundTruth_v.txt', delimiter=' ')          # corner format -> xywh (:36-41)
        x_min, y_min = np.min(c[:, [0, 2]], axis=1)[:, None], np.min(c[:, [1, 3]], axis=1)[:, None]
        x_max, y_max = np.max(c[:, [0, 2]], axis=1)[:, None], np.max(c[:, [1, 3]], axis=1)[:, None]
        gt = np.concatenate((x_min, y_min, x_max - x_min, y_max - y_min), axis=1)
    elif 'VTUAV' in set_type:
        rgb, x = imgs('rgb', '.jpg'), imgs('ir', '.jpg')
        gt = np.loadtxt(seq_path + '/rgb.txt', delimiter=' ')
    else:
        raise ValueError("unknown RGBT set_type %r" % (set_type,))
    return rgb, x, np.atleast_2d(gt)


def gen_config_rgbe(seq_path, set_type):
    """-> (rgb paths, event paths, gt, absent_label), trimmed to the first present frame
    (test_rgbe_mgpus.py:18-27,59-63)."""
    if set_type not in ('VisEvent', 'visevent'):
        raise ValueError("unknown RGBE set_type %r" % (set_type,))
    rgb = sorted([seq_path + '/vis_imgs/' + p for p in os.listdir(seq_path + '/vis_imgs') if os.path.splitext(p)[1] == '.bmp'])
    ev = sorted([seq_path + '/event_imgs/' + p for p in os.listdir(seq_path + '/event_imgs') if os.path.splitext(p)[1] == '.bmp'])
    gt = np.atleast_2d(np.loadtxt(seq_path + '/groundtruth.txt', delimiter=','))
    absent = np.atleast_1d(np.loadtxt(seq_path + '/absent_label.txt'))
    if absent[0] == 0:
        first = int(absent.argmax())
        rgb, ev, gt = rgb[first:], ev[first:], gt[first:]
    return rgb, ev, gt, absent


def new_result(n_frames, gt):
    """result array of the reference loops: zeros_like(gt) when the lengths agree, row 0 = ground truth
    (test_rgbt_mgpus.py:93-97)."""
    result = np.zeros_like(gt) if n_frames == len(gt) else np.zeros((n_frames, 4), dtype=gt.dtype)
    result[0] = np.copy(gt[0])
    return result


def save_result(path, result, kind):
    """Byte-compatible writers: RGBT `np.savetxt(path, result)` (test_rgbt_mgpus.py:116), RGBE
    `np.savetxt(path, result, fmt='%.14f', delimiter=',')` (test_rgbe_mgpus.py:83)."""
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    if kind == 'rgbt':
        np.savetxt(path, result)
    elif kind == 'rgbe':
        np.savetxt(path, result, fmt='%.14f', delimiter=',')
    else:
        raise ValueError(kind)


def plan_groups(frame_counts, batch):
    """Sequences sorted by length (longest first) and cut into groups of `batch`, so that the sequences of a
    group finish close together.  -> list of lists of indices."""
    order = sorted(range(len(frame_counts)), key=lambda i: (-int(frame_counts[i]), i))
    return [order[i:i + batch] for i in range(0, len(order), batch)]


def run_sequences(seq_names, seq_home, dataset_name, state_dict, save_dir, kind='rgbt', prompt_type='vipt_deep',
                  batch=8, xtype='rgbrgb', rank=0, world=1, skip_existing=True, load_frame=None):
    """Track every sequence of `seq_names` assigned to this rank and write `<save_dir>/<seq>.txt`.
    Returns {sequence name: result (T,4)}.  `load_frame(rgb_path, x_path, dtype)` defaults to
    vipt_b200.frames.get_x_frame (host decode, device assembly)."""
    from .frames import get_x_frame
    from .tracker import BatchedViPTTracker
    load_frame = load_frame or (lambda a, b, dtype: get_x_frame(a, b, dtype=dtype))
    cfgs = {}
    todo = []
    for name in seq_names:
        seq_txt = name.split('/')[1] if 'VTUAV' in dataset_name else name          # test_rgbt_mgpus.py:67-70
        path = os.path.join(save_dir, seq_txt + '.txt')
        if skip_existing and os.path.exists(path):
            continue                                                                # :77-79
        cfg = gen_config_rgbt(seq_home + '/' + name, dataset_name) if kind == 'rgbt' else \
            gen_config_rgbe(seq_home + '/' + name, dataset_name)[:3]
        cfgs[name] = (cfg, path)
        todo.append(name)
    # like `zip(RGB_img_list, T_img_list)` (test_rgbt_mgpus.py:98): a sequence is as long as its shorter image list
    nfr = {n: min(len(cfgs[n][0][0]), len(cfgs[n][0][1])) for n in todo}
    mine = [todo[i] for i in shard.assign_sequences([nfr[n] for n in todo], world)[rank]] if todo else []
    results = {}
    failed = {}
    trackers = {}

    def tracker_for(nb):
        if nb not in trackers:                       # the engine is built for a fixed number of slots
            trackers[nb] = BatchedViPTTracker(state_dict, nb, prompt_type, use_graph=True)
        return trackers[nb]

    def small_box(e):
        return "Too small bounding box" in str(e)

    for group in plan_groups([nfr[n] for n in mine], batch):
        names = [mine[i] for i in group]
        # initialize; a sequence whose first box is degenerate ('Too small bounding box.', processing_utils.py:34) is
        # dropped -- in the reference only that sequence's process dies -- and the rest of the group starts again
        while names:
            trk = tracker_for(len(names))
            lists = [cfgs[n][0] for n in names]
            frames = [load_frame(l[0][0], l[1][0], xtype) for l in lists]
            try:
                trk.initialize(frames, [l[2][0].tolist() for l in lists])
                break
            except Exception as e:
                if not small_box(e):
                    raise
                bad = trk.failed_slots()
                for k in bad:
                    failed[names[k]] = "Too small bounding box. (frame 0)"
                    print("vipt_b200.driver: %s: Too small bounding box. (frame 0) -- sequence skipped" % names[k])
                names = [n for k, n in enumerate(names) if k not in bad]
        if not names:
            continue
        res = [new_result(nfr[n], cfgs[n][0][2]) for n in names]
        alive = [True] * len(names)
        T = max(nfr[n] for n in names)
        for t in range(1, T):
            # a finished (or failed) sequence keeps its last frame; its boxes are no longer recorded
            frames = [load_frame(l[0][t], l[1][t], xtype) if (t < nfr[n] and alive[k]) else frames[k]
                      for k, (n, l) in enumerate(zip(names, lists))]
            try:
                boxes, _ = trk.track(frames)
            except Exception as e:
                if not small_box(e):
                    raise
                for k in trk.failed_slots():          # only the offending sequences are lost (their slots idle on)
                    if alive[k]:
                        alive[k] = False
                        failed[names[k]] = "Too small bounding box. (frame %d)" % t
                        print("vipt_b200.driver: %s: Too small bounding box. (frame %d) -- sequence dropped" % (names[k], t))
                boxes, _ = trk.collect()
            for k, n in enumerate(names):
                if t < nfr[n] and alive[k]:
                    res[k][t] = boxes[k]
        for k, (n, r) in enumerate(zip(names, res)):
            if alive[k]:
                save_result(cfgs[n][1], r, kind)
                results[n] = r
    run_sequences.failed = failed
    return results
