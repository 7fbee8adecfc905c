"""Run the UNMODIFIED reference (wxltop/Multi-Modal-Trakcing-Bechmark, ViPT/) on CPU -- TEST INFRASTRUCTURE ONLY.

This file only works inside the build container, where /root/reference exists.  Nothing in
`-m gpu` tests, smoke() or bench.py imports it; it is used by `tests/golden/make_golden.py`
(to write the committed golden vectors) and by the `not gpu` tests that pin `oracle/vipt_oracle.py`
against the real reference when the checkout is present.

What it does:
  * puts `oracle/stubs` (timm 0.5.4 subset, easydict, vot, visdom, jpeg4py, matplotlib) ahead of
    the reference on sys.path,
  * patches `.cuda()` to identity (ViPT/lib/test/tracker/vipt.py:23,30 and data_utils.py:17-22
    hard-code `.cuda()`),
  * builds `TrackerParams` by hand (ViPT/lib/test/parameter/vipt.py:7-29 reads the author's
    absolute prj_dir) and a checkpoint file {'net': state_dict}.
"""
import os
import sys
import tempfile

import torch

# the checkout in the build container, else the byte-for-byte copy of the path's files that oracle/make_ref.py put
# under oracle/_ref (git-ignored; travels to the GPU box with the snapshot)
_TRAVEL = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "ViPT")
REF_ROOT = os.environ.get("VIPT_REFERENCE_ROOT") or (
    "/root/reference/ViPT" if os.path.isdir("/root/reference/ViPT/lib") else _TRAVEL)
_STUBS = os.path.join(os.path.dirname(os.path.abspath(__file__)), "stubs")


def available() -> bool:
    return os.path.isdir(os.path.join(REF_ROOT, "lib", "models", "vipt"))


_imported = False


def import_reference():
    global _imported
    if _imported:
        return
    if not available():
        raise RuntimeError("reference checkout not present at %s" % REF_ROOT)
    for p in (REF_ROOT, _STUBS):
        if p in sys.path:
            sys.path.remove(p)
    sys.path.insert(0, REF_ROOT)
    sys.path.insert(0, _STUBS)
    # .cuda() -> identity so the reference's hard-coded device moves run on CPU
    torch.Tensor.cuda = lambda self, *a, **k: self
    torch.nn.Module.cuda = lambda self, *a, **k: self
    _imported = True


def make_cfg(yaml_name: str):
    """cfg exactly as ViPT/lib/test/parameter/vipt.py:12-14 builds it."""
    import_reference()
    import importlib
    config = importlib.import_module("lib.config.vipt.config")
    config = importlib.reload(config)  # fresh defaults for every call
    yaml_file = os.path.join(REF_ROOT, "experiments", "vipt", yaml_name + ".yaml")
    config.update_config_from_file(yaml_file)
    return config.cfg


def build_reference_model(yaml_name: str):
    import_reference()
    from lib.models.vipt import build_viptrack
    cfg = make_cfg(yaml_name)
    model = build_viptrack(cfg, training=False)
    model.eval()
    return model, cfg


def make_tracker(yaml_name: str, state_dict):
    """Unmodified ViPTTrack (ViPT/lib/test/tracker/vipt.py:17) over a synthetic checkpoint."""
    import_reference()
    from lib.test.tracker.vipt import ViPTTrack
    from lib.test.utils import TrackerParams
    cfg = make_cfg(yaml_name)
    params = TrackerParams()
    params.cfg = cfg
    params.template_factor = cfg.TEST.TEMPLATE_FACTOR
    params.template_size = cfg.TEST.TEMPLATE_SIZE
    params.search_factor = cfg.TEST.SEARCH_FACTOR
    params.search_size = cfg.TEST.SEARCH_SIZE
    params.save_all_boxes = False
    params.debug = 0
    fd, path = tempfile.mkstemp(suffix=".pth")
    os.close(fd)
    try:
        torch.save({"net": state_dict}, path)
        params.checkpoint = path
        tracker = ViPTTrack(params)
    finally:
        os.unlink(path)
    return tracker
