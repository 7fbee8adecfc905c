"""ctypes binding of libvipt_b200.so (include/vipt_b200.h).  No fallback: a missing library is fatal."""
import ctypes
import os

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_PKG, "libvipt_b200.so")
_EXPERIMENT = os.environ.get("VIPT_EXPERIMENT_LIB")     # kernel experiments only (scripts/): another build of the library

_lib = None

c_void_p = ctypes.c_void_p
c_int = ctypes.c_int
c_i64 = ctypes.c_int64
c_double = ctypes.c_double
c_float = ctypes.c_float

# name -> (restype, argtypes); must list every symbol include/vipt_b200.h declares
SIGNATURES = {
    "vipt_version": (ctypes.c_char_p, []),
    "vipt_last_error": (ctypes.c_char_p, []),
    "vipt_op_gemm": (c_int, [c_void_p, c_i64, c_void_p, c_i64, c_int, c_int, c_int, c_int, c_int, c_void_p, c_i64,
                             c_void_p, c_void_p, c_void_p]),
    "vipt_op_gemm_resid_stat": (c_int, [c_void_p, c_i64, c_void_p, c_i64, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p,
                                        c_void_p, c_void_p]),
    "vipt_op_gemm_lnfold": (c_int, [c_void_p, c_i64, c_void_p, c_i64, c_int, c_int, c_int, c_int, c_void_p, c_i64, c_void_p,
                                    c_void_p, c_void_p, c_int, c_float, c_void_p, c_void_p, c_void_p]),
    "vipt_op_gemm_resid_prompt": (c_int, [c_void_p, c_i64, c_void_p, c_i64, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p,
                                          c_void_p, c_void_p, c_void_p]),
    "vipt_op_crop": (c_int, [c_void_p, c_void_p, c_void_p, c_double, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p,
                             c_int, c_void_p]),
    "vipt_op_attention": (c_int, [c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "vipt_op_ce_topk": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                c_int, c_void_p]),
    "vipt_op_conv3x3": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_i64,
                                c_i64, c_void_p, c_void_p]),
    "vipt_op_assemble_rgbx": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p]),
    "vipt_op_assemble_rgbd": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "vipt_op_assemble_rgbd_noalloc": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p]),
    "vipt_engine_create": (c_int, [ctypes.POINTER(c_void_p)]),
    "vipt_engine_destroy": (c_int, [c_void_p]),
    "vipt_engine_set_tensor": (c_int, [c_void_p, ctypes.c_char_p, c_void_p, c_i64]),
    "vipt_engine_finalize": (c_int, [c_void_p, c_int, c_int]),
    "vipt_engine_forward": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p]),
    "vipt_engine_init": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "vipt_engine_track_enqueue": (c_int, [c_void_p, c_void_p, c_int, c_void_p]),
    "vipt_engine_slot_errors": (c_int, [c_void_p, c_void_p, c_void_p]),
    "vipt_engine_init_host": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "vipt_engine_track_host": (c_int, [c_void_p, c_void_p, c_int, ctypes.POINTER(c_i64), c_void_p]),
    "vipt_engine_collect": (c_int, [c_void_p, c_void_p, c_void_p, ctypes.POINTER(c_i64)]),
    "vipt_engine_set_state": (c_int, [c_void_p, c_void_p, c_void_p]),
    "vipt_engine_profile_step": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "vipt_engine_results": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p]),
    "vipt_engine_info": (c_i64, [c_void_p, ctypes.c_char_p]),
    "vipt_engine_buffer": (c_int, [c_void_p, ctypes.c_char_p, ctypes.POINTER(c_void_p), ctypes.POINTER(c_i64)]),
}


class ViptError(RuntimeError):
    pass


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ViptError("libvipt_b200.so is not built (%s); run `python __graft_entry__.py build` -- "
                            "there is no CPU fallback" % LIB_PATH)
        L = ctypes.CDLL(_EXPERIMENT if _EXPERIMENT else LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(code, what=""):
    if code != 0:
        msg = lib().vipt_last_error().decode()
        if code == -5:
            raise Exception("Too small bounding box.")   # same text as processing_utils.py:35
        raise ViptError("%s failed with code %d: %s" % (what, code, msg))


def ptr(t):
    """device pointer of a torch tensor (or None)."""
    return None if t is None else c_void_p(t.data_ptr())


def stream_ptr():
    import torch
    return c_void_p(torch.cuda.current_stream().cuda_stream)
