"""VOT / TraX consumer of the tracker: the loop of ViPT/lib/test/vot/vipt_class.py:19-113.

    from vipt_b200 import vot_driver
    vot_driver.run_vot_exp(params, handle, out_conf=True)

`handle` is the reference's `vot.VOT("rectangle", channels=...)` object (ViPT/lib/test/vot/vot.py; it needs the
vot-toolkit / trax packages, which are not in this image): anything with `region()`, `frame()` and
`report(region, confidence=None)`.  `frame()` returns one image path, or `[color_path, depth_path]` for the rgbd
channels, or a false value at the end of the sequence -- exactly what the reference loop consumes.  Per frame:

    get_rgbd_frame(color, depth, dtype='rgbcolormap', depth_clip=True)   (vipt_class.py:77-79)  device assembly
    tracker.initialize(image, {'init_bbox': [x, y, w, h]})               (:33-41)
    tracker.track(image) -> (target_bbox, best_score)                    (:43-48)
    handle.report(Rectangle(*bbox), score)                               (:95-98)

The tracker is ViPTTrack in batch-1 latency mode (CUDA-graph replay).  The reference's `vis` branch
(cv2.rectangle / imwrite of every frame, :99-113) is a debugging aid and is not reproduced.
"""
import collections

from . import frames as _frames
from .tracker import ViPTTrack

Rectangle = collections.namedtuple("Rectangle", ["x", "y", "width", "height"])      # vot.Rectangle (vot.py)


class vipt(object):
    """Same three-method object as vipt_class.py:19-48, over a ready `params` (ViPT/lib/test/parameter/vipt.py)."""

    def __init__(self, params):
        params.visualization = False
        params.debug = False
        self.tracker = ViPTTrack(params)

    def initialize(self, img_rgb, selection):
        x, y, w, h = selection
        self.H, self.W = int(img_rgb.shape[0]), int(img_rgb.shape[1])
        self.tracker.initialize(img_rgb, {'init_bbox': [x, y, w, h]})

    def track(self, img_rgb):
        outputs = self.tracker.track(img_rgb)
        return outputs['target_bbox'], outputs['best_score']


def _read(imagefile):
    """vipt_class.py:76-80 / :89-93: an rgbd pair goes through get_rgbd_frame, a single file is decoded BGR -> RGB."""
    if isinstance(imagefile, (list, tuple)) and len(imagefile) == 2:
        return _frames.get_rgbd_frame(imagefile[0], imagefile[1], dtype='rgbcolormap', depth_clip=True)
    raise NotImplementedError("single-image (RGB only) VOT channels: ViPT is a two-modality tracker; the reference "
                              "loop feeds its 3-channel image to a 6-channel model and fails in PatchEmbed")


def run_vot_exp(params, handle, out_conf=False, rectangle=Rectangle):
    """The loop of vipt_class.py:51-98.  Returns the number of frames tracked (the reference returns nothing)."""
    tracker = vipt(params)
    selection = handle.region()
    imagefile = handle.frame()
    if not imagefile:
        return 0
    tracker.initialize(_read(imagefile), selection)
    n = 0
    while True:
        imagefile = handle.frame()
        if not imagefile:
            break
        b1, max_score = tracker.track(_read(imagefile))
        if out_conf:
            handle.report(rectangle(*b1), max_score)
        else:
            handle.report(rectangle(*b1))
        n += 1
    return n
