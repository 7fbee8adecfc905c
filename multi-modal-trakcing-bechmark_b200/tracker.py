"""Drop-in for the reference tracker API (ViPT/lib/test/tracker/vipt.py:17-130):

    tracker = ViPTTrack(params)            # params.cfg, params.checkpoint, template/search factor + size
    tracker.initialize(image, {'init_bbox': [x, y, w, h]})
    out = tracker.track(image)             # {'target_bbox': [x, y, w, h], 'best_score': float}

`image` is the reference's np.uint8 (H, W, 6) RGB||X frame.  Everything between the frame and the box
(crop, resize, normalise, backbone, head, Hann window, argmax, box decode, map_box_back, clip_box) runs
in libvipt_b200.so; the box state lives on the device in fp64.  `BatchedViPTTracker` is the same loop
for many independent sequences at once (what the reference does with one process per sequence,
ViPT/RGBT_workspace/test_rgbt_mgpus.py:66-117).
"""
import numpy as np
import torch

from .engine import NativeEngine
from .model import _cfg_get


def _load_state_dict(params):
    sd = getattr(params, "state_dict", None)
    if sd is None:
        sd = torch.load(params.checkpoint, map_location="cpu")["net"]      # vipt.py:21
    return sd


class BatchedViPTTracker:
    """B independent sequences advanced in lock-step on one GPU."""

    def __init__(self, state_dict, batch, prompt_type="vipt_deep", use_graph=True):
        self.engine = NativeEngine(state_dict, batch, prompt_type)
        self.batch = batch
        self.use_graph = use_graph
        self.frame_id = 0
        self._dev_frames = None
        self._ptrs = None
        self._host_mode = False
        self.h2d_bytes = 0

    @staticmethod
    def _is_host(im):
        return isinstance(im, np.ndarray) or (isinstance(im, torch.Tensor) and not im.is_cuda)

    @staticmethod
    def _host_view(im):
        """-> (address, (H, W), object to keep alive) of a host uint8 (H, W, 6) frame."""
        if isinstance(im, np.ndarray):
            a = np.ascontiguousarray(im)
            assert a.dtype == np.uint8 and a.ndim == 3 and a.shape[2] == 6, "frames are uint8 (H, W, 6)"
            return a.ctypes.data, a.shape[:2], a
        t = im.contiguous()
        assert t.dtype == torch.uint8 and t.dim() == 3 and t.shape[2] == 6, "frames are uint8 (H, W, 6)"
        return t.data_ptr(), tuple(t.shape[:2]), t

    def initialize(self, images, init_boxes):
        boxes = np.asarray(init_boxes, np.float64).reshape(self.batch, 4)
        self._host_mode = all(self._is_host(im) for im in images)
        if self._host_mode:
            views = [self._host_view(im) for im in images]
            self._hw = [v[1] for v in views]
            self.engine.init_sequences_host([v[0] for v in views], self._hw, boxes, keepalive=[v[2] for v in views])
        else:
            frames = [im if (isinstance(im, torch.Tensor) and im.is_cuda) else
                      torch.from_numpy(np.ascontiguousarray(im)).to(self.engine.device) for im in images]
            self._order_after_producer(frames)
            self.engine.init_sequences(frames, boxes)
        self.frame_id = 0
        self.state = boxes.copy()
        self.h2d_bytes = 0
        self._pending_host = 0

    def _order_after_producer(self, frames):
        """Device frames are usually produced on the caller's current stream (frames.get_x_frame, .to(device)); the
        engine runs on its own stream, so it must wait for the producer, and the caching allocator must not hand the
        frame memory out again before the engine's crop kernel has read it."""
        es = self.engine.stream
        es.wait_stream(torch.cuda.current_stream(self.engine.device))
        for f in frames:
            f.record_stream(es)

    def failed_slots(self):
        """Indices of the sequences that raised 'Too small bounding box.' since the last initialize (their results are
        meaningless from that frame on; the other sequences are unaffected).  Lowers the engine's error flag."""
        return [int(i) for i in np.nonzero(self.engine.slot_errors())[0]]

    def set_state(self, boxes):
        """Override the box state of every sequence (teacher-forced evaluation)."""
        self.engine.set_state(boxes)
        self.state = np.asarray(boxes, np.float64).reshape(self.batch, 4).copy()

    def track_async(self, images):
        if self._host_mode and all(self._is_host(im) for im in images):
            views = [self._host_view(im) for im in images]
            assert [v[1] for v in views] == self._hw, "frame size changed within a sequence"
            # only the rectangle the search crop reads travels to the device (csrc/engine3.cu)
            n = self.engine.track_host([v[0] for v in views], self.use_graph, keepalive=[v[2] for v in views])
            if n >= 0:
                self.h2d_bytes = n
            self._pending_host += 1
        else:
            frames = [im if (isinstance(im, torch.Tensor) and im.is_cuda) else
                      torch.from_numpy(np.ascontiguousarray(im)).to(self.engine.device) for im in images]
            self._order_after_producer(frames)
            with torch.cuda.stream(self.engine.stream):
                ptrs = torch.tensor([f.data_ptr() for f in frames], dtype=torch.int64).pin_memory().to(
                    self.engine.device, non_blocking=True)
            self._dev_frames, self._ptrs = frames, ptrs        # keep alive until the step has run
            self.engine.track_enqueue(ptrs, self.use_graph)
        self.frame_id += 1

    def collect(self):
        """Boxes (B,4) and scores (B,) of the oldest step enqueued with track_async(host frames) and not yet collected.
        With page-locked host frames up to 4 steps may be in flight: enqueue step t+1, then collect step t, so the
        host never stalls the device (the rectangle of step t+1 is evaluated on the device from step t's boxes)."""
        if self._pending_host > 0:
            self._pending_host -= 1          # the step leaves the ring even when it reports 'Too small bounding box.'
            boxes, scores, n = self.engine.collect()
            self.h2d_bytes = n
        else:
            boxes, scores = self.engine.results()
        self.state = boxes
        return boxes, scores

    def track(self, images):
        self.track_async(images)
        while self._pending_host > 1:            # results of earlier track_async calls nobody collected
            self.collect()
        return self.collect()


class ViPTTrack:
    """Single-sequence mirror of the reference class (batch-1 latency mode, CUDA-graph replay)."""

    def __init__(self, params):
        self.params = params
        self.cfg = params.cfg
        prompt_type = _cfg_get(self.cfg, "TRAIN.PROMPT.TYPE", "vipt_deep")
        for name, want in (("template_size", 128), ("search_size", 256), ("template_factor", 2.0),
                           ("search_factor", 4.0)):
            got = getattr(params, name, want)
            if float(got) != float(want):
                raise NotImplementedError("params.%s = %r (engine is built for %r)" % (name, got, want))
        self._impl = BatchedViPTTracker(_load_state_dict(params), 1, prompt_type)
        self.state = None
        self.frame_id = 0
        self.debug = getattr(params, "debug", 0) or 0
        self.save_all_boxes = getattr(params, "save_all_boxes", False)

    def initialize(self, image, info: dict):
        self._impl.initialize([image], [info['init_bbox']])
        self.state = list(info['init_bbox'])
        self.frame_id = 0
        if self.save_all_boxes:
            return {"all_boxes": list(info['init_bbox'])}

    def track(self, image, info: dict = None):
        self.frame_id += 1
        boxes, scores = self._impl.track([image])
        self.state = [float(v) for v in boxes[0]]
        out = {"target_bbox": self.state, "best_score": float(scores[0])}
        if self.save_all_boxes:
            out["all_boxes"] = list(self.state)
        return out


def get_tracker_class():
    return ViPTTrack
