"""CPU restatement of the reference's crop / resize / normalise step -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this.  The product path never does.

Follows (reference file:line, relative to /root/reference/):
  * sample_target                 ViPT/lib/train/data/processing_utils.py:14-81
  * PreprocessorMM.process        ViPT/lib/test/tracker/data_utils.py:15-24
  * cv.resize (INTER_LINEAR, u8)  third-party dependency opencv-python (4.13 in this image; the
    reference pins none): published algorithm restated here -- 11-bit fixed-point coefficients,
    horizontal pass in int32, vertical pass `(((b0*(r0>>4))>>16) + ((b1*(r1>>4))>>16) + 2) >> 2`,
    and the exact-2x downscale special case that OpenCV routes to INTER_AREA.
Pinned against cv2 itself in tests/test_oracle_crop.py (cv2 is importable in the image).
"""
import math

import numpy as np

INTER_BITS = 11
INTER_SCALE = 1 << INTER_BITS


def crop_geometry(box, factor, im_h, im_w):
    """processing_utils.py:27-47.  Returns (x1, y1, crop_sz) in Python ints (round = half-to-even)."""
    x, y, w, h = box
    crop_sz = math.ceil(math.sqrt(w * h) * factor)
    if crop_sz < 1:
        raise Exception('Too small bounding box.')
    x1 = round(x + 0.5 * w - crop_sz * 0.5)
    y1 = round(y + 0.5 * h - crop_sz * 0.5)
    return x1, y1, crop_sz


def crop_padded(im, x1, y1, crop_sz):
    """processing_utils.py:36-55: slice + zero border (cv.BORDER_CONSTANT).  Note the `+1` in the
    right/bottom pad: the last image row/column is never sampled when the crop overhangs."""
    H, W = im.shape[:2]
    x2, y2 = x1 + crop_sz, y1 + crop_sz
    x1_pad = max(0, -x1); x2_pad = max(x2 - W + 1, 0)
    y1_pad = max(0, -y1); y2_pad = max(y2 - H + 1, 0)
    im_crop = im[y1 + y1_pad:y2 - y2_pad, x1 + x1_pad:x2 - x2_pad, :]
    out = np.zeros((im_crop.shape[0] + y1_pad + y2_pad, im_crop.shape[1] + x1_pad + x2_pad, im.shape[2]), np.uint8)
    out[y1_pad:y1_pad + im_crop.shape[0], x1_pad:x1_pad + im_crop.shape[1]] = im_crop
    return out


def _linear_coeffs(src, dst, clamp_frac=True):
    """OpenCV resize(): scale = 1/(dst/src) in double; fx = float((d+0.5)*scale-0.5).
    Horizontal taps zero the fraction at the borders (clamp_frac=True); vertical taps keep the
    fraction and only clip the two row indices (resizeGeneric_Invoker), which rounds differently."""
    inv_scale = float(dst) / float(src)
    scale = 1.0 / inv_scale
    idx = np.empty(dst, np.int64)
    coef = np.empty((dst, 2), np.int32)
    for d in range(dst):
        fx = np.float32((d + 0.5) * scale - 0.5)
        sx = int(math.floor(float(fx)))
        fx = np.float32(fx - np.float32(sx))
        if clamp_frac:
            if sx < 0:
                fx = np.float32(0); sx = 0
            if sx >= src - 1:
                fx = np.float32(0); sx = src - 1
        c0 = np.float32(np.float32(1.0) - fx) * np.float32(INTER_SCALE)
        c1 = np.float32(fx) * np.float32(INTER_SCALE)
        idx[d] = sx
        coef[d, 0] = int(np.rint(c0))   # saturate_cast<short>(float) = cvRound = round-half-even
        coef[d, 1] = int(np.rint(c1))
    return idx, coef


def resize_bilinear_u8(src, out_sz):
    """cv.resize(src, (out_sz, out_sz)) default INTER_LINEAR on uint8 HxWxC, restated."""
    sh, sw = src.shape[:2]
    if sh == 2 * out_sz and sw == 2 * out_sz:
        # resize(): INTER_LINEAR with integer scale 2x2 is routed to INTER_AREA.  With 6 channels
        # OpenCV has no fast-area kernel and takes the generic float path: saturate_cast(sum*0.25)
        # = round-half-to-EVEN (1-3-4 channel images would get (sum+2)>>2 instead).
        s = src.astype(np.int32)
        acc = s[0::2, 0::2] + s[0::2, 1::2] + s[1::2, 0::2] + s[1::2, 1::2]
        q = acc >> 2; r = acc & 3
        return (q + ((r == 3) | ((r == 2) & ((q & 1) == 1)))).astype(np.uint8)
    xi, xa = _linear_coeffs(sw, out_sz)
    yi, ya = _linear_coeffs(sh, out_sz, clamp_frac=False)
    s = src.astype(np.int32)
    xi1 = np.minimum(xi + 1, sw - 1)   # when xi == sw-1 the second coefficient is 0
    # horizontal pass: int32 rows scaled by 2^11
    rows = s[:, xi, :] * xa[None, :, 0, None] + s[:, xi1, :] * xa[None, :, 1, None]
    yi0 = np.clip(yi, 0, sh - 1); yi1 = np.clip(yi + 1, 0, sh - 1)
    r0 = rows[yi0]; r1 = rows[yi1]
    b0 = ya[:, 0][:, None, None]; b1 = ya[:, 1][:, None, None]
    out = (((b0 * (r0 >> 4)) >> 16) + ((b1 * (r1 >> 4)) >> 16) + 2) >> 2
    return np.clip(out, 0, 255).astype(np.uint8)


def sample_target(im, target_bb, search_area_factor, output_sz, use_cv2=False):
    """processing_utils.py:14-76 (att_mask omitted: ViPTTrack never uses it, vipt.py:43,67).
    use_cv2=True runs the reference's own third-party call for the resize (baseline timing)."""
    x1, y1, crop_sz = crop_geometry(target_bb, search_area_factor, im.shape[0], im.shape[1])
    padded = crop_padded(im, x1, y1, crop_sz)
    resize_factor = output_sz / crop_sz
    if use_cv2:
        import cv2
        return cv2.resize(padded, (output_sz, output_sz)), resize_factor
    return resize_bilinear_u8(padded, output_sz), resize_factor


MEAN6 = np.array([0.485, 0.456, 0.406, 0.485, 0.456, 0.406], np.float32)
STD6 = np.array([0.229, 0.224, 0.225, 0.229, 0.224, 0.225], np.float32)


def preprocess(patch_u8):
    """data_utils.py:20-24: (x/255 - mean)/std in fp32, HWC -> (1,6,S,S)."""
    x = patch_u8.astype(np.float32).transpose(2, 0, 1)[None]
    return ((x / np.float32(255.0)) - MEAN6.reshape(1, 6, 1, 1)) / STD6.reshape(1, 6, 1, 1)
