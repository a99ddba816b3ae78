"""ctypes binding of libb200vq.so (include/b200vq.h).  Fails loudly: there is no fallback.

The shared object is built in-tree by ``video_tokenizer_b200.build`` (nvcc, sm_100a) and
loaded from the package directory, so the process's loaded-library list shows it.
"""
from __future__ import annotations

import ctypes as C
import os
import threading

from . import build as _build

_lock = threading.Lock()
_lib = None

DT_F32, DT_BF16, DT_F16 = 0, 1, 2
MODE_ARGMIN_DIST, MODE_ARGMAX_COS, MODE_SAMPLE = 0, 1, 2

_i64, _i32, _f32, _u64, _vp, _sz = C.c_int64, C.c_int, C.c_float, C.c_uint64, C.c_void_p, C.c_size_t

# name -> (restype, argtypes); must list every symbol include/b200vq.h declares
SIGNATURES = {
    "b200vq_version": (_i32, []),
    "b200vq_error_string": (C.c_char_p, [_i32]),
    "b200vq_last_cuda_error": (_i32, []),
    "b200vq_max_dim": (_i32, []),
    "b200vq_launch_count": (_u64, []),
    "b200vq_fma_probe": (_i32, [_i64, _vp, _vp, _vp]),
    "b200vq_workspace_bytes": (_sz, [_i64, _i32, _i64]),
    "b200vq_prepare_codebook": (_i32, [_vp, _i64, _i32, _i32, _vp, _vp, _vp, _vp]),
    "b200vq_keys_reset": (_i32, [_vp, _i64, _vp]),
    "b200vq_scores_partial": (_i32, [_vp, _i32, _i64, _i64, _i32, _vp, _vp, _i64, _i64, _i32, _i32, _f32,
                                     _u64, _u64, _i64, _vp, _vp]),
    "b200vq_finish": (_i32, [_vp, _i32, _i64, _i64, _i32, _vp, _i64, _vp, _vp, _i32, _i64, _vp, _vp, _vp, _vp,
                             _vp, _vp, _vp, _i32, _i64, _f32, _f32, _vp, _vp]),
    "b200vq_tc_supported": (_i32, [_i32, _i32, _i32]),
    "b200vq_tc_image_bytes": (_sz, [_i64, _i32]),
    "b200vq_tc_scratch_bytes": (_sz, [_i64]),
    "b200vq_prepare_codebook_tc": (_i32, [_vp, _i64, _i32, _vp, _vp]),
    "b200vq_scores_tc": (_i32, [_vp, _i32, _i64, _i64, _i32, _vp, _vp, _vp, _i64, _i64, _i32, _i32, _f32, _u64, _u64,
                                _i64, _vp, _vp, _vp, _vp]),
    "b200vq_forward": (_i32, [_vp, _i32, _i64, _i64, _i32, _vp, _vp, _vp, _i64, _i32, _i32, _f32, _u64, _u64,
                              _i64, _f32, _f32, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "b200vq_backward": (_i32, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i64, _i64, _i32, _i64, _i32, _i64, _f32, _f32,
                               _vp, _i32, _vp, _vp]),
    "b200vq_codebook_backward": (_i32, [_vp, _vp, _vp, _vp, _i64, _i32, _i32, _vp, _vp]),
    "b200vq_decode": (_i32, [_vp, _i64, _vp, _i64, _i32, _i32, _vp, _vp, _vp]),
    "b200vq_usage_stats": (_i32, [_vp, _vp, _i64, _vp, _vp]),
    "b200vq_linear_supported": (_i32, [_i32, _i32]),
    "b200vq_linear_in": (_i32, [_vp, _i64, _i64, _i32, _vp, _vp, _i32, _vp, _vp]),
    "b200vq_linear_out": (_i32, [_vp, _i64, _i32, _vp, _i32, _vp, _i32, _vp, _i64, _vp]),
    "b200vq_weight_grad": (_i32, [_vp, _i64, _vp, _i64, _i32, _i32, _vp, _i32, _vp, _vp]),
    "b200vq_decode_project": (_i32, [_vp, _i64, _vp, _i64, _i32, _i32, _vp, _vp, _i32, _vp, _i64, _vp, _i64, _vp, _vp]),
    "b200vq_pack3": (_i32, [_vp, _vp, _vp, _vp, _vp]),
    "b200vq_entropy_supported": (_i32, [_i32]),
    "b200vq_entropy_scratch_bytes": (_sz, [_i64, _i64]),
    "b200vq_entropy_forward": (_i32, [_vp, _i64, _i32, _vp, _i64, _f32, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "b200vq_entropy_backward": (_i32, [_vp, _i64, _i32, _vp, _i64, _f32, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "b200vq_grad_out_supported": (_i32, [_i32, _i32]),
    "b200vq_grad_out": (_i32, [_vp, _i64, _vp, _vp, _i64, _i32, _i32, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "b200vq_weight_grad_workspace_bytes": (_sz, [_i32, _i32]),
    "b200vq_weight_grad_ws": (_i32, [_vp, _i64, _vp, _i64, _i32, _i32, _vp, _i32, _vp, _vp, _sz, _vp]),
    "b200vq_weight_grad_noalloc": (_i32, [_vp, _i64, _vp, _i64, _i32, _i32, _vp, _i32, _vp, _vp]),
    "b200vq_colsum": (_i32, [_vp, _i64, _i32, _vp, _vp]),
    "b200vq_gather_owned": (_i32, [_vp, _i64, _vp, _i64, _i64, _i32, _vp, _vp]),
    "b200vq_bottleneck_workspace_bytes": (_sz, [_i64, _i32, _i32, _i64]),
    "b200vq_bottleneck_forward": (_i32, [_vp, _i64, _i64, _i32, _vp, _vp, _vp, _i64, _i32, _vp, _vp, _i32, _i32, _f32,
                                         _u64, _u64, _i64, _f32, _f32, _i32] + [_vp] * 13 + [_i64, _vp, _sz, _vp, _i32, _vp, _i64, _vp]),
    "b200vq_bottleneck_backward": (_i32, [_vp, _i32, _i64, _vp, _vp, _vp, _vp, _vp, _vp, _i64, _i64, _i32] + [_vp] * 8
                                   + [_i64, _i32, _i32, _f32, _f32, _vp, _i64] + [_vp] * 6 + [_sz, _vp, _vp]),
}


class B200VQError(RuntimeError):
    pass


def lib_path() -> str:
    return _build.LIB


def load(build_if_missing: bool = True):
    """Load (building first when stale and nvcc is present) and type every export."""
    global _lib
    with _lock:
        if _lib is not None:
            return _lib
        path = lib_path()
        if build_if_missing and _build.is_stale():
            try:
                _build.build()
            except Exception as e:  # stale-but-present library is still usable on the GPU box
                if not os.path.exists(path):
                    raise B200VQError(f"libb200vq.so is missing and could not be built: {e}") from e
        if not os.path.exists(path):
            raise B200VQError(f"{path} not found: run `python -m video_tokenizer_b200.build` (no CPU fallback exists)")
        lib = C.CDLL(path)
        for name, (res, args) in SIGNATURES.items():
            try:
                fn = getattr(lib, name)
            except AttributeError as e:
                raise B200VQError(f"{path} does not export {name}") from e
            fn.restype, fn.argtypes = res, args
        if lib.b200vq_version() < 122:
            raise B200VQError("libb200vq.so is older than this Python package")
        _lib = lib
        return lib


def check(rc: int, what: str):
    if rc == 0:
        return
    lib = load()
    msg = lib.b200vq_error_string(rc).decode()
    if rc == -4:
        msg += f" [cudaError {lib.b200vq_last_cuda_error()}]"
    if rc in (-1, -2, -3):
        raise ValueError(f"{what}: {msg}")
    raise B200VQError(f"{what}: {msg}")
