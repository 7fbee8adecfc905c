"""Recipe for oracle/_ref: the UNMODIFIED reference files of the tracking path, made to travel to the GPU box.

    python oracle/make_ref.py          (build container only: needs /root/reference)

The reference is Python, so there is nothing to compile: the recipe imports the reference's tracker once
(oracle/ref_harness.py), collects every module that import pulled in from /root/reference/ViPT (45 files, the
closure of lib/test/tracker/vipt.py) plus the experiment yaml files, and copies them byte for byte to
oracle/_ref/ViPT/.  oracle/_ref/ is git-ignored (no reference source enters the history) but not gpurun-ignored,
so on the GPU box `bench.py --impl reference` and the CPU-baseline leg can run the reference's own classes
(`cpu_baseline.kind = "reference"`).  __graft_entry__.build() calls this when /root/reference is present.
"""
import glob
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
SRC = "/root/reference/ViPT"
DST = os.path.join(HERE, "_ref", "ViPT")


def make(force=False):
    if not os.path.isdir(os.path.join(SRC, "lib", "models", "vipt")):
        return None
    stamp = os.path.join(DST, ".complete")
    if os.path.exists(stamp) and not force:
        return DST
    sys.path.insert(0, ROOT)
    os.environ["VIPT_REFERENCE_ROOT"] = SRC
    from oracle import ref_harness as rh
    from vipt_b200 import synth
    rh.REF_ROOT = SRC
    rh.make_tracker("deep_rgbt", synth.synthetic_state_dict("vipt_deep", 0))
    files = sorted(set(m.__file__ for m in list(sys.modules.values())
                       if getattr(m, "__file__", None) and m.__file__.startswith(SRC + os.sep)))
    files += glob.glob(os.path.join(SRC, "experiments", "vipt", "*.yaml"))
    for f in files:
        d = os.path.join(DST, os.path.relpath(f, SRC))
        os.makedirs(os.path.dirname(d), exist_ok=True)
        shutil.copyfile(f, d)
    open(stamp, "w").write("%d files\n" % len(files))
    return DST


if __name__ == "__main__":
    print(make(force="--force" in sys.argv))
