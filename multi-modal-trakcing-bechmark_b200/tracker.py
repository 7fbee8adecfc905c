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

    def _upload(self, images):
        """host uint8 frames -> device (pinned staging, async copy); returns device tensors."""
        out = []
        with torch.cuda.stream(self.engine.stream):
            for im in images:
                if isinstance(im, torch.Tensor) and im.is_cuda:
                    out.append(im)
                else:
                    t = torch.from_numpy(np.ascontiguousarray(im)) if not isinstance(im, torch.Tensor) else im
                    assert t.dtype == torch.uint8 and t.dim() == 3 and t.shape[2] == 6, "frames are uint8 (H, W, 6)"
                    if not t.is_pinned():
                        t = t.pin_memory()
                    out.append(t.to(self.engine.device, non_blocking=True))
        return out

    def initialize(self, images, init_boxes):
        frames = self._upload(images)
        self.engine.init_sequences(frames, np.asarray(init_boxes, np.float64))
        self.frame_id = 0
        self.state = np.asarray(init_boxes, np.float64).reshape(self.batch, 4).copy()

    def track_async(self, images):
        frames = self._upload(images)
        with torch.cuda.stream(self.engine.stream):
            ptrs = torch.tensor([f.data_ptr() for f in frames], dtype=torch.int64).pin_memory().to(
                self.engine.device, non_blocking=True)
        self._dev_frames, self._ptrs = frames, ptrs        # keep alive until the step has run
        self.engine.track_enqueue(ptrs, self.use_graph)
        self.frame_id += 1

    def track(self, images):
        self.track_async(images)
        boxes, scores = self.engine.results()
        self.state = boxes
        return boxes, scores


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
