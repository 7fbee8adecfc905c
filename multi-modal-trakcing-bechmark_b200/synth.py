"""Synthetic weights and sequences.

There is no network for checkpoints or datasets, so both the reference (oracle) side and the B200
engine are fed the same seeded random-init `state_dict` and the same seeded synthetic frames.

`synthetic_state_dict` produces exactly the key set / shapes that the reference's
`ViPTTrack.__init__` loads with `strict=True` (ViPT/lib/test/tracker/vipt.py:21; key list from
`build_viptrack(cfg).state_dict()`): 352 tensors for `vipt_deep`, 253 for `vipt_shaw`.
The distributions follow the reference's default init (ViPT/lib/models/vipt/vit.py:191-223,
ViPT/lib/models/layers/head.py:126-128, ViPT/lib/models/vipt/vit_ce_prompt.py:58-60) with the
adjustments SURVEY.md 8d recommends so that the parity test is discriminative (non-zero position
embeddings, non-trivial LN / BN statistics, a size bias that keeps the box stable, a peaked score map).
"""
import math

import numpy as np
import torch

EMBED = 768
DEPTH = 12
MLP = 3072
HIDE = 8
HEAD_CH = 256


def _tn(g, shape, std):
    t = torch.empty(shape)
    torch.nn.init.trunc_normal_(t, mean=0.0, std=std, a=-2.0, b=2.0, generator=g)
    return t


def _xavier(g, shape):
    fan_out = shape[0] * int(np.prod(shape[2:]))
    fan_in = shape[1] * int(np.prod(shape[2:]))
    a = math.sqrt(6.0 / (fan_in + fan_out))
    return (torch.rand(shape, generator=g) * 2 - 1) * a


# Named weight sets.  "flat": the reference's default init (qkv std 0.02): attention is almost uniform, so the
# candidate-elimination scores of all search tokens sit within a few 1e-4 of 1/N of each other and most frames are
# decided by differences below fp16-operand resolution.  "peaked": qkv std 0.06 gives attention logits of std ~2.9
# (max ~15; row entropy ~4.0 nats against 5.8 for uniform attention, CE scores spread over 2.5 decades), the range a
# trained ViT works in.  Larger values are not usable as a parity workload: the random network turns chaotic (logit
# std 11 at qkv std 0.12: near-argmax attention, where rounding the GEMM operands of the fp32 reference itself to
# 16 bits changes the CE scores by ~100 % and the score map by ~50 %; measured with oracle.vipt_oracle.OPERAND_ROUND).
WEIGHT_SETS = {"flat": 0.02, "peaked": 0.06}


def weight_set(name: str, prompt_type: str = "vipt_deep", seed: int = 0):
    """state_dict of a named weight set (WEIGHT_SETS); "outlier" = the peaked set with outlier channels (below)."""
    if name == "outlier":
        return with_outlier_channels(synthetic_state_dict(prompt_type, seed, qkv_std=WEIGHT_SETS["peaked"]))
    return synthetic_state_dict(prompt_type, seed, qkv_std=WEIGHT_SETS[name])


def with_outlier_channels(sd, scale=40.0, channels=(5, 301, 640), blocks=(1, 4, 7, 11)):
    """Trained ViTs carry a few channels whose LayerNorm gain is tens of times the typical one (and matching large
    fc1 weights), which is what stresses 16-bit GEMM operands: the LayerNorm outputs of those channels reach
    ~scale * 3, the fc1 pre-activations and the GELU outputs (fp16 in the engine) grow accordingly.  In place."""
    for i in blocks:
        p = "backbone.blocks.%d." % i
        for c in channels:
            sd[p + "norm1.weight"][c] *= scale
            sd[p + "norm2.weight"][c] *= scale
            sd[p + "mlp.fc1.weight"][:, c] *= 4.0
    return sd


def synthetic_state_dict(prompt_type: str = "vipt_deep", seed: int = 0, qkv_std: float = 0.02):
    """name -> fp32 tensor, in the reference's state_dict order."""
    assert prompt_type in ("vipt_deep", "vipt_shaw")
    g = torch.Generator().manual_seed(seed)
    sd = {}
    n = lambda *s: torch.randn(*s, generator=g)
    bb = "backbone."
    sd[bb + "cls_token"] = _tn(g, (1, 1, EMBED), 0.02)           # dead weight (vit.py:137)
    sd[bb + "pos_embed"] = _tn(g, (1, 197, EMBED), 0.02)         # dead weight (vit.py:139)
    sd[bb + "pos_embed_z"] = _tn(g, (1, 64, EMBED), 0.02)
    sd[bb + "pos_embed_x"] = _tn(g, (1, 256, EMBED), 0.02)
    for pe in ("patch_embed", "patch_embed_prompt"):
        k = 1.0 / math.sqrt(3 * 16 * 16)
        sd[bb + pe + ".proj.weight"] = (torch.rand((EMBED, 3, 16, 16), generator=g) * 2 - 1) * k
        sd[bb + pe + ".proj.bias"] = (torch.rand((EMBED,), generator=g) * 2 - 1) * k
    n_prompt = DEPTH if prompt_type == "vipt_deep" else 1
    for i in range(n_prompt):
        p = bb + "prompt_blocks.%d." % i
        sd[p + "conv0_0.weight"] = _xavier(g, (HIDE, EMBED, 1, 1))
        sd[p + "conv0_0.bias"] = 0.05 * n(HIDE)
        sd[p + "conv0_1.weight"] = _xavier(g, (HIDE, EMBED, 1, 1))
        sd[p + "conv0_1.bias"] = 0.05 * n(HIDE)
        sd[p + "conv1x1.weight"] = _xavier(g, (EMBED, HIDE, 1, 1))
        sd[p + "conv1x1.bias"] = 0.05 * n(EMBED)
        sd[p + "fovea.smooth"] = torch.full((1,), 10.0)
    for i in range(n_prompt):
        p = bb + "prompt_norms.%d." % i
        sd[p + "weight"] = 1.0 + 0.1 * n(EMBED)
        sd[p + "bias"] = 0.1 * n(EMBED)
    for i in range(DEPTH):
        p = bb + "blocks.%d." % i
        sd[p + "norm1.weight"] = 1.0 + 0.1 * n(EMBED)
        sd[p + "norm1.bias"] = 0.1 * n(EMBED)
        sd[p + "attn.qkv.weight"] = _tn(g, (3 * EMBED, EMBED), qkv_std)
        sd[p + "attn.qkv.bias"] = 0.02 * n(3 * EMBED)
        sd[p + "attn.proj.weight"] = _tn(g, (EMBED, EMBED), 0.02)
        sd[p + "attn.proj.bias"] = 0.02 * n(EMBED)
        sd[p + "norm2.weight"] = 1.0 + 0.1 * n(EMBED)
        sd[p + "norm2.bias"] = 0.1 * n(EMBED)
        sd[p + "mlp.fc1.weight"] = _tn(g, (MLP, EMBED), 0.02)
        sd[p + "mlp.fc1.bias"] = 0.02 * n(MLP)
        sd[p + "mlp.fc2.weight"] = _tn(g, (EMBED, MLP), 0.02)
        sd[p + "mlp.fc2.bias"] = 0.02 * n(EMBED)
    sd[bb + "norm.weight"] = 1.0 + 0.1 * n(EMBED)
    sd[bb + "norm.bias"] = 0.1 * n(EMBED)
    chans = [EMBED, HEAD_CH, HEAD_CH // 2, HEAD_CH // 4, HEAD_CH // 8]
    for tower, nout in (("ctr", 1), ("offset", 2), ("size", 2)):
        for li in range(4):
            p = "box_head.conv%d_%s." % (li + 1, tower)
            cin, cout = chans[li], chans[li + 1]
            sd[p + "0.weight"] = _xavier(g, (cout, cin, 3, 3))
            sd[p + "0.bias"] = 0.05 * n(cout)
            sd[p + "1.weight"] = 1.0 + 0.1 * n(cout)
            sd[p + "1.bias"] = 0.1 * n(cout)
            sd[p + "1.running_mean"] = 0.1 * n(cout)
            sd[p + "1.running_var"] = 0.5 + torch.rand(cout, generator=g)
            sd[p + "1.num_batches_tracked"] = torch.tensor(1, dtype=torch.int64)
        p = "box_head.conv5_%s." % tower
        w = _xavier(g, (nout, chans[4], 1, 1))
        b = 0.05 * n(nout)
        if tower == "ctr":
            w = w * 5.0                                       # score-map contrast without saturating the clamp
        if tower == "size":
            b = torch.full((nout,), math.log(0.25 / 0.75))    # sigmoid -> 0.25: stable box size
        sd[p + "weight"] = w
        sd[p + "bias"] = b
    return sd


def reorder_like(sd, ref_keys):
    """Return sd re-ordered to the reference's key order (strict load does not need it; goldens do)."""
    return {k: sd[k] for k in ref_keys}


DATASET_SHAPES = {  # (H, W) builder's assumption, SURVEY.md 8d
    "rgbd": (360, 640),    # DepthTrack-shaped
    "rgbt": (480, 630),    # LasHeR-shaped
    "rgbe": (260, 346),    # VisEvent-shaped
}


def _smooth_noise(rng, H, W, C, cell=16):
    """Low-pass-filtered noise: bilinear upsample of a coarse random grid."""
    gh, gw = H // cell + 2, W // cell + 2
    coarse = rng.random((gh, gw, C), dtype=np.float32)
    ys = np.arange(H, dtype=np.float32) / cell
    xs = np.arange(W, dtype=np.float32) / cell
    y0 = ys.astype(np.int32); x0 = xs.astype(np.int32)
    fy = (ys - y0)[:, None, None]; fx = (xs - x0)[None, :, None]
    a = coarse[y0][:, x0]; b = coarse[y0][:, x0 + 1]
    c = coarse[y0 + 1][:, x0]; d = coarse[y0 + 1][:, x0 + 1]
    return (a * (1 - fy) * (1 - fx) + b * (1 - fy) * fx + c * fy * (1 - fx) + d * fy * fx)


def synthetic_sequence(seed: int, n_frames: int, modality: str = "rgbt", box_wh=(60, 48)):
    """frames: uint8 (T, H, W, 6) RGB||X; init_box [x, y, w, h] (floats).

    Smooth random background + a textured rectangle on a random walk; the second modality is a
    different random colour map of the same scene."""
    H, W = DATASET_SHAPES[modality]
    rng = np.random.default_rng(seed)
    bg = _smooth_noise(rng, H, W, 3)
    mix = rng.random((3, 3), dtype=np.float32)
    mix /= mix.sum(0, keepdims=True)
    bw, bh = box_wh
    tex = rng.random((bh, bw, 3), dtype=np.float32)
    tex = 0.5 * tex + 0.5 * _smooth_noise(rng, bh, bw, 3, cell=8)
    cx = W * (0.35 + 0.3 * rng.random()); cy = H * (0.35 + 0.3 * rng.random())
    frames = np.empty((n_frames, H, W, 6), dtype=np.uint8)
    boxes = np.empty((n_frames, 4), dtype=np.float64)
    vx, vy = 0.0, 0.0
    for t in range(n_frames):
        if t > 0:
            vx = 0.8 * vx + 1.5 * rng.standard_normal(); vy = 0.8 * vy + 1.5 * rng.standard_normal()
            cx = float(np.clip(cx + vx, bw, W - bw)); cy = float(np.clip(cy + vy, bh, H - bh))
        x0 = int(round(cx - bw / 2)); y0 = int(round(cy - bh / 2))
        img = bg.copy()
        img[y0:y0 + bh, x0:x0 + bw] = tex
        img += 0.02 * rng.standard_normal(img.shape).astype(np.float32)
        aux = img @ mix
        frames[t, :, :, :3] = np.clip(img * 255.0, 0, 255).astype(np.uint8)
        frames[t, :, :, 3:] = np.clip(aux * 255.0, 0, 255).astype(np.uint8)
        boxes[t] = (x0, y0, bw, bh)
    return frames, [float(v) for v in boxes[0]], boxes


_KEYS_CACHE = {}


def synthetic_state_dict_keys(prompt_type):
    """Key set of the reference state_dict (used for strict loading)."""
    if prompt_type not in _KEYS_CACHE:
        keys = []
        bb = "backbone."
        keys += [bb + k for k in ("cls_token", "pos_embed", "pos_embed_z", "pos_embed_x")]
        for pe in ("patch_embed", "patch_embed_prompt"):
            keys += [bb + pe + ".proj.weight", bb + pe + ".proj.bias"]
        n_prompt = DEPTH if prompt_type == "vipt_deep" else 1
        for i in range(n_prompt):
            p = bb + "prompt_blocks.%d." % i
            keys += [p + s for s in ("conv0_0.weight", "conv0_0.bias", "conv0_1.weight", "conv0_1.bias",
                                     "conv1x1.weight", "conv1x1.bias", "fovea.smooth")]
        for i in range(n_prompt):
            keys += [bb + "prompt_norms.%d.weight" % i, bb + "prompt_norms.%d.bias" % i]
        for i in range(DEPTH):
            p = bb + "blocks.%d." % i
            for m in ("norm1", "attn.qkv", "attn.proj", "norm2", "mlp.fc1", "mlp.fc2"):
                keys += [p + m + ".weight", p + m + ".bias"]
        keys += [bb + "norm.weight", bb + "norm.bias"]
        for tower in ("ctr", "offset", "size"):
            for li in range(1, 5):
                p = "box_head.conv%d_%s." % (li, tower)
                keys += [p + s for s in ("0.weight", "0.bias", "1.weight", "1.bias", "1.running_mean",
                                         "1.running_var", "1.num_batches_tracked")]
            keys += ["box_head.conv5_%s.weight" % tower, "box_head.conv5_%s.bias" % tower]
        _KEYS_CACHE[prompt_type] = keys
    return _KEYS_CACHE[prompt_type]
