"""Thin Python owner of a native engine (libvipt_b200.so).  torch is used only for device memory,
streams and host<->device copies; every computation happens in the CUDA library."""
import collections
import ctypes

import numpy as np
import torch

from . import _lib

_DTYPES = {
    "score_map": (torch.float32, (-1, 1, 16, 16)), "size_map": (torch.float32, (-1, 2, 16, 16)),
    "offset_map": (torch.float32, (-1, 2, 16, 16)), "pred_boxes": (torch.float32, (-1, 4)),
    "removed": (torch.int32, (3, -1, 256)), "ce_score": (torch.float32, (3, -1, 256)),
    "feat32": (torch.float32, (-1, 320, 768)), "feat16": (torch.float16, (-1, 256, 768)),
    "x": (torch.float32, None), "xA": (torch.float32, None), "xB": (torch.float32, None), "xn": (torch.float16, None), "dbuf": (torch.float16, None),
    "FV": (torch.float32, (-1, 320, 8)), "XD": (torch.float32, (-1, 320, 16)), "E": (torch.float32, (2, -1, 320, 768)),
    "qkv": (torch.float16, None), "att": (torch.float16, None), "hbuf": (torch.float16, None),
    "gi": (torch.int32, (-1, 256)), "slot_of_pos": (torch.int32, (-1, 256)), "perm": (torch.int32, (-1, 256)),
    "h1": (torch.float16, (-1, 256, 768)), "h4": (torch.float16, (-1, 256, 96)),
    "state": (torch.float64, (-1, 4)), "out_boxes": (torch.float64, (-1, 4)), "out_scores": (torch.float32, (-1,)),
    "geom_x": (torch.float64, (-1, 4)), "geom_z": (torch.float64, (-1, 4)), "apatch": (torch.float16, (2, -1, 320, 768)),
    "ce_prob": (torch.float32, (-1, 12, 256)),
}


class _RawBuffer:
    """__cuda_array_interface__ view of engine-owned device memory."""

    def __init__(self, ptr, nbytes, owner):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}
        self._owner = owner


class NativeEngine:
    def __init__(self, state_dict, batch, prompt_type="vipt_deep", device=None):
        if not torch.cuda.is_available():
            raise _lib.ViptError("vipt_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
        if prompt_type not in ("vipt_deep", "vipt_shaw"):
            raise ValueError("unsupported TRAIN.PROMPT.TYPE %r" % (prompt_type,))
        self.L = _lib.lib()
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.batch = int(batch)
        self.prompt_type = prompt_type
        h = ctypes.c_void_p()
        # the engine runs on its own stream: the legacy default stream cannot be graph-captured
        self.stream = torch.cuda.Stream(device=self.device)
        with torch.cuda.device(self.device):
            _lib.check(self.L.vipt_engine_create(ctypes.byref(h)), "vipt_engine_create")
            self.h = h
            for name, t in state_dict.items():
                if not t.dtype.is_floating_point:
                    continue                      # num_batches_tracked
                a = np.ascontiguousarray(t.detach().cpu().float().numpy())
                _lib.check(self.L.vipt_engine_set_tensor(self.h, name.encode(), a.ctypes.data_as(ctypes.c_void_p), a.size),
                           "vipt_engine_set_tensor(%s)" % name)
            _lib.check(self.L.vipt_engine_finalize(self.h, self.batch, 1 if prompt_type == "vipt_deep" else 0),
                       "vipt_engine_finalize")
        self._frames_keepalive = None
        # host frames of every enqueued, not yet collected track_host step: the gather kernel and the speculative
        # uploads read them asynchronously (csrc/engine3.cu), so they must outlive the call that enqueued them
        self._host_keepalive = collections.deque()

    def __del__(self):
        try:
            if getattr(self, "h", None):
                self.L.vipt_engine_destroy(self.h)
                self.h = None
        except Exception:
            pass

    # ------------------------------------------------------------------ buffers
    def buffer(self, name):
        ptr = ctypes.c_void_p()
        nbytes = ctypes.c_int64()
        _lib.check(self.L.vipt_engine_buffer(self.h, name.encode(), ctypes.byref(ptr), ctypes.byref(nbytes)),
                   "vipt_engine_buffer(%s)" % name)
        raw = torch.as_tensor(_RawBuffer(ptr.value, nbytes.value, self), device=self.device)
        dtype, shape = _DTYPES[name]
        t = raw.view(dtype)
        return t.view(*shape) if shape is not None else t

    def info(self, what):
        return int(self.L.vipt_engine_info(self.h, what.encode()))

    # ------------------------------------------------------------------ model-level forward
    def forward(self, template, search, want_feat=False, debug_stop=0):
        assert template.shape == (self.batch, 6, 128, 128) and search.shape == (self.batch, 6, 256, 256), \
            "engine was built for batch %d" % self.batch
        template = template.to(self.device, torch.float32).contiguous()
        search = search.to(self.device, torch.float32).contiguous()
        self.stream.wait_stream(torch.cuda.current_stream(self.device))
        _lib.check(self.L.vipt_engine_forward(self.h, _lib.ptr(template), _lib.ptr(search), int(want_feat), int(debug_stop),
                                              self._sp()), "vipt_engine_forward")
        self.stream.synchronize()      # outputs are read by the caller's stream

    def _sp(self):
        return ctypes.c_void_p(self.stream.cuda_stream)

    # ------------------------------------------------------------------ tracking
    def init_sequences(self, frames, boxes):
        """frames: list of B uint8 CUDA tensors (H,W,6); boxes: (B,4) [x,y,w,h]."""
        assert len(frames) == self.batch
        self._frames_keepalive = frames
        ptrs = (ctypes.c_void_p * self.batch)(*[f.data_ptr() for f in frames])
        hw = np.ascontiguousarray(np.array([[f.shape[0], f.shape[1]] for f in frames], np.int32))
        bx = np.ascontiguousarray(np.asarray(boxes, np.float64).reshape(self.batch, 4))
        _lib.check(self.L.vipt_engine_init(self.h, ptrs, hw.ctypes.data_as(ctypes.c_void_p),
                                           bx.ctypes.data_as(ctypes.c_void_p), self._sp()), "vipt_engine_init")

    def init_sequences_host(self, host_ptrs, hw, boxes, keepalive=None):
        """host_ptrs: B host addresses of uint8 (H,W,6) frames; hw (B,2) int32; boxes (B,4)."""
        self._frames_keepalive = keepalive
        self._host_keepalive.clear()
        ptrs = (ctypes.c_void_p * self.batch)(*host_ptrs)
        hw = np.ascontiguousarray(np.asarray(hw, np.int32).reshape(self.batch, 2))
        bx = np.ascontiguousarray(np.asarray(boxes, np.float64).reshape(self.batch, 4))
        _lib.check(self.L.vipt_engine_init_host(self.h, ptrs, hw.ctypes.data_as(ctypes.c_void_p),
                                                bx.ctypes.data_as(ctypes.c_void_p), self._sp()), "vipt_engine_init_host")

    def track_host(self, host_ptrs, use_graph=True, keepalive=None):
        """Upload the search rectangles of B host frames and enqueue one step; returns bytes uploaded.
        The frames must stay valid and unmodified until this step has been collected (include/vipt_b200.h); the
        `keepalive` objects are held until then."""
        self._host_keepalive.append(keepalive)
        ptrs = (ctypes.c_void_p * self.batch)(*host_ptrs)
        n = ctypes.c_int64(0)
        _lib.check(self.L.vipt_engine_track_host(self.h, ptrs, int(use_graph), ctypes.byref(n), self._sp()),
                   "vipt_engine_track_host")
        return n.value

    def collect(self):
        """-> (boxes (B,4) f64, scores (B,) f32, bytes uploaded) of the oldest uncollected track_host step."""
        boxes = np.empty((self.batch, 4), np.float64)
        scores = np.empty((self.batch,), np.float32)
        n = ctypes.c_int64(0)
        _lib.check(self.L.vipt_engine_collect(self.h, boxes.ctypes.data_as(ctypes.c_void_p),
                                              scores.ctypes.data_as(ctypes.c_void_p), ctypes.byref(n)), "vipt_engine_collect")
        if self._host_keepalive:
            self._host_keepalive.popleft()
        return boxes, scores, n.value

    def slot_errors(self):
        """(B,) int32: sticky per-sequence error flags (0 or -5 'Too small bounding box.'); lowers the engine-wide flag."""
        out = np.zeros(self.batch, np.int32)
        _lib.check(self.L.vipt_engine_slot_errors(self.h, out.ctypes.data_as(ctypes.c_void_p), self._sp()),
                   "vipt_engine_slot_errors")
        return out

    def set_state(self, boxes):
        bx = np.ascontiguousarray(np.asarray(boxes, np.float64).reshape(self.batch, 4))
        _lib.check(self.L.vipt_engine_set_state(self.h, bx.ctypes.data_as(ctypes.c_void_p), self._sp()),
                   "vipt_engine_set_state")

    def track_enqueue(self, frame_ptrs_dev, use_graph=True):
        """frame_ptrs_dev: int64 CUDA tensor of B frame device pointers.  Asynchronous."""
        _lib.check(self.L.vipt_engine_track_enqueue(self.h, _lib.ptr(frame_ptrs_dev), int(use_graph), self._sp()),
                   "vipt_engine_track_enqueue")

    KERNEL_CLASSES = ("gemm_linear", "gemm_head", "attention", "prompt", "norm_ce", "crop", "im2col", "head_tail")

    def profile_step(self, frame_ptrs_dev):
        """{class: (ms, launches)} for one un-graphed step, CUDA events around every launch."""
        ms = np.zeros(8, np.float32)
        n = np.zeros(8, np.int32)
        _lib.check(self.L.vipt_engine_profile_step(self.h, _lib.ptr(frame_ptrs_dev), ms.ctypes.data_as(ctypes.c_void_p),
                                                   n.ctypes.data_as(ctypes.c_void_p), self._sp()),
                   "vipt_engine_profile_step")
        return {k: (float(ms[i]), int(n[i])) for i, k in enumerate(self.KERNEL_CLASSES)}

    def results(self):
        boxes = np.empty((self.batch, 4), np.float64)
        scores = np.empty((self.batch,), np.float32)
        _lib.check(self.L.vipt_engine_results(self.h, boxes.ctypes.data_as(ctypes.c_void_p),
                                              scores.ctypes.data_as(ctypes.c_void_p), self._sp()),
                   "vipt_engine_results")
        self._host_keepalive.clear()        # the stream was synchronised: every enqueued step has read its frames
        return boxes, scores
