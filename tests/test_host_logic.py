"""CPU-side checks: the C ABI loads and exports every declared symbol (no compute calls without a GPU),
host-side sharding / gather over gloo with world_size 2, config handling of the reference-facing mirrors."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from vipt_b200 import _lib, build, shard

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_c_abi_exports_every_declared_symbol():
    build.build()
    hdr = open(os.path.join(ROOT, "include", "vipt_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(vipt_[a-z_0-9]+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)
    L = ctypes.CDLL(_lib.LIB_PATH)
    for name in declared:
        assert getattr(L, name) is not None
    assert b"sm_100a" in _lib.lib().vipt_version()


def test_library_has_blackwell_sass():
    """the hot kernels are tcgen05 / TMA code, not a recompiled legacy path"""
    import shutil
    import subprocess
    if shutil.which("cuobjdump") is None:
        pytest.skip("cuobjdump not on PATH")
    sass = subprocess.run(["cuobjdump", "-sass", _lib.LIB_PATH], capture_output=True, text=True).stdout
    assert "UTCHMMA" in sass and "UTMALDG" in sass and "LDTM" in sass
    assert "HMMA.16" not in sass        # no mma.sync tensor path


def test_assign_sequences_balanced_and_deterministic():
    counts = [50, 10, 10, 10, 40, 30, 20, 5]
    a = shard.assign_sequences(counts, 2)
    assert sorted(a[0] + a[1]) == list(range(8))
    loads = [sum(counts[i] for i in r) for r in a]
    assert abs(loads[0] - loads[1]) <= 10
    assert a == shard.assign_sequences(counts, 2)
    assert shard.assign_sequences([7, 7, 7], 1) == [[0, 1, 2]]
    assert shard.assign_sequences([], 4) == [[], [], [], []]


def _worker(rank, world, port, counts, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    mine = shard.assign_sequences(counts, world)[rank]
    boxes = [np.arange(counts[i] * 4, dtype=np.float64).reshape(-1, 4) + 1000 * i for i in mine]
    out = shard.gather_boxes(mine, boxes, len(counts))
    ok = all(np.array_equal(out[i], np.arange(counts[i] * 4, dtype=np.float64).reshape(-1, 4) + 1000 * i)
             for i in range(len(counts)))
    q.put((rank, ok, sorted(out)))
    dist.destroy_process_group()


def test_gather_boxes_gloo_world2():
    counts = [5, 3, 8, 1, 4]          # ragged sequence lengths
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 500
    ps = [ctx.Process(target=_worker, args=(r, 2, port, counts, q)) for r in range(2)]
    [p.start() for p in ps]
    res = [q.get(timeout=120) for _ in range(2)]
    [p.join(30) for p in ps]
    assert all(ok for _, ok, _ in res), res
    assert all(ids == list(range(5)) for _, _, ids in res)


def test_model_mirror_config_handling_without_gpu():
    from vipt_b200.model import build_viptrack
    m = build_viptrack({"TRAIN": {"PROMPT": {"TYPE": "vipt_shaw"}}})
    assert m.prompt_type == "vipt_shaw"
    with pytest.raises(NotImplementedError):
        build_viptrack({"DATA": {"SEARCH": {"SIZE": 320}}})
    with pytest.raises(NotImplementedError):
        build_viptrack({"MODEL": {"BACKBONE": {"TYPE": "vit_base_patch16_224_prompt"}}})
    if not torch.cuda.is_available():
        from vipt_b200.engine import NativeEngine
        with pytest.raises(_lib.ViptError, match="no CPU fallback"):
            NativeEngine({}, 1)
