"""ctypes binding of libavtubes.so (the C ABI declared in include/avtubes.h).

There is no CPU fallback: if the library is missing or the device is not sm_100 every
entry point raises.  PyTorch only supplies device memory and the current stream.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
# AVT_LIB_PATH: development only -- load an experimental build (scripts/build_variant.sh) instead
LIB_PATH = os.environ.get("AVT_LIB_PATH") or os.path.join(HERE, "libavtubes.so")

PRECISION_FP32 = 0
PRECISION_F16 = 1
GT_VGGSS, GT_FLICKR, GT_FLICKR_FRAME = 0, 1, 2
EVAL_WORKSPACE_BYTES = 64
MTC_STATIC_MAX_FRAMES = 4096
DTYPE_F32, DTYPE_F16, DTYPE_BF16 = 0, 1, 2
STAGE_PREP, STAGE_SIM, BWD_ROWGRAD_DONE = 1, 2, 1

_p = C.c_void_p
_i = C.c_int
_f = C.c_float
_z = C.c_size_t
_d = C.c_double
_ll = C.c_longlong

# name -> (restype, argtypes); mirrors include/avtubes.h line for line
SIGNATURES = {
    "avt_version": (_i, []),
    "avt_last_error": (C.c_char_p, []),
    "avt_device_ok": (_i, []),
    "avt_launch_count": (C.c_longlong, []),
    "avt_profile_enable": (_i, [_i]),
    "avt_profile_read": (_i, [_i, _p, _p]),
    "avt_audio_normalize": (_i, [_p, _i, _i, _p, _p, _p]),
    "avt_audio_normalize_backward": (_i, [_p, _p, _p, _i, _i, _p, _p]),
    "avt_head_workspace_bytes": (_z, [_i, _i, _i, _i, _i, _i]),
    "avt_head_forward": (_i, [_p, _i, _p, _i, _i, _i, _i, _p, _i, _f, _f, _f, _f, _f, _i, _i, _i, _i,
                              _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _i, _f, _p, _z, _p]),
    "avt_head_backward": (_i, [_p, _i, _p, _i, _i, _i, _i, _p, _i, _f, _f, _f, _f, _f, _i, _i, _i, _i,
                               _p, _p, _p, _p, _p, _p, _f, _p, _p, _p, _p, _p, _p, _i, _p, _p, _z, _p]),
    "avt_eval_maps": (_i, [_p, _i, _p, _p, _p, _i, _i, _p, _p, _p, _p, _p, _p, _p]),
    "avt_eval_stats": (_i, [_p, _i]),
    "avt_ciou_counts": (_i, [_p, _f, _p, _i, _i, _p, _p, _p, _p]),
    "avt_rasterise_gt": (_i, [_p, _p, _i, _i, _i, _p, _p]),
    "avt_auc": (_i, [_p, _i, _p, _p, _p]),
    "avt_mtc": (_i, [_p, _i, _p, _p, _p, _p]),
    "avt_mtc_static": (_i, [_p, _i, _p, _p]),
    "avt_spectrogram_frames": (_i, [_i]),
    "avt_log_spectrogram": (_i, [_p, _p, _i, _ll, _i, _p, _d, _d, _d, _p, _i, _p]),
    "avt_log_spectrogram_pcm16": (_i, [_p, _p, _i, _ll, _i, _p, _d, _d, _d, _p, _i, _p]),
    "avt_video_metrics": (_i, [_p, _p, _i, _i, _p, _p, _p, _p]),
}

_lib = None


def load():
    """dlopen the in-tree library (once) and attach the prototypes."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -m audio_visual_tubes_b200.build` "
            "(there is no CPU or PyTorch fallback for this path)")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)           # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def last_error() -> str:
    return load().avt_last_error().decode("utf-8", "replace")


def check(rc: int, what: str):
    if rc != 0:
        raise RuntimeError(f"{what} failed (status {rc}): {last_error()}")


def ptr(t):
    """Device pointer of a torch tensor (None -> NULL)."""
    return None if t is None else C.c_void_p(t.data_ptr())


def stream_ptr():
    import torch
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def require_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise RuntimeError("libavtubes runs on the GPU only: got a CPU tensor "
                               "(there is no CPU fallback; move inputs to cuda)")


PROF_SLOTS = {"prep": 0, "sim": 1, "finish": 2, "row_grad": 3, "bwd_g": 4, "bwd_dv": 5,
              "bwd_da": 6, "dv_finish": 7, "eval": 8}


def profile_enable(on: bool = True):
    check(load().avt_profile_enable(int(on)), "avt_profile_enable")


def profile_read():
    """{stage: (total_ms, intervals)} since the last profile_enable()."""
    lib = load()
    out = {}
    for name, slot in PROF_SLOTS.items():
        ms, n = C.c_double(0.0), C.c_int(0)
        check(lib.avt_profile_read(slot, C.byref(ms), C.byref(n)), "avt_profile_read")
        out[name] = (ms.value, n.value)
    return out


def launch_count() -> int:
    return int(load().avt_launch_count())
