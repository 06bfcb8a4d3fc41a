"""ctypes binding of ``libfedml_b200.so`` (C ABI in include/fedml_b200.h).

The library is built in-tree by :mod:`fedml_b200.native.build` and loaded once.  On a box with
a GPU a missing/unloadable library is a hard error — there is deliberately no silent PyTorch
fallback for the device path (the eager implementations in ``fedml_b200.core.aggregation`` are
CPU oracles, selected explicitly by the caller when tensors live on the CPU).
"""
from __future__ import annotations

import ctypes as C
import os
import threading

import torch

_LIB = None
_LOCK = threading.Lock()
LIB_PATH = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "native", "libfedml_b200.so")

c_f32p, c_u32p, c_i32p, c_u64p, c_i64p, c_vp = (C.c_void_p,) * 6  # raw device pointers travel as integers

FB_OPT = {"none": 0, "fedavg": 0, "sgd": 1, "adam": 2, "fedadam": 2, "adagrad": 3, "fedadagrad": 3, "yogi": 4,
          "fedyogi": 4}
FB_ROUND_MAX_BLOCKS = 1024


class FbRoundParams(C.Structure):
    _fields_ = [
        ("clients", C.c_void_p), ("ld", C.c_int64), ("C", C.c_int32), ("client_base", C.c_int32),
        ("weights", C.c_void_p), ("weights_buf", C.c_void_p), ("P", C.c_int64), ("P_train", C.c_int64),
        ("inv_total", C.c_float), ("inv_total_buf", C.c_float), ("self_coef", C.c_float),
        ("global_model", C.c_void_p), ("partial", C.c_void_p), ("peer_partial", C.c_void_p),
        ("peer_global", C.c_void_p), ("peer_flags", C.c_void_p), ("rank", C.c_int32), ("G", C.c_int32),
        ("epoch", C.c_uint32), ("two_shot", C.c_int32), ("clip_bound", C.c_float), ("noise_std", C.c_float),
        ("seed", C.c_uint64), ("round", C.c_uint32), ("opt", C.c_int32), ("nesterov", C.c_int32),
        ("lr", C.c_float), ("momentum", C.c_float), ("beta1", C.c_float), ("beta2", C.c_float), ("eps", C.c_float),
        ("bias_corr1", C.c_float), ("bias_corr2", C.c_float), ("grad_scale", C.c_float),
        ("m", C.c_void_p), ("v", C.c_void_p), ("sqnorm_part", C.c_void_p), ("clip_scale_out", C.c_void_p),
        ("grid_bar", C.c_void_p), ("grid_bar_target", C.c_uint32),
        ("mc_partial", C.c_void_p), ("mc_global", C.c_void_p), ("agg_noise_std", C.c_float),
    ]


class FbGroupedGemm(C.Structure):
    _fields_ = [
        ("a", C.c_void_p), ("b", C.c_void_p), ("d", C.c_void_p), ("bias", C.c_void_p),
        ("a_gs", C.c_int64), ("b_gs", C.c_int64), ("d_gs", C.c_int64), ("bias_gs", C.c_int64),
        ("G", C.c_int32), ("M", C.c_int32), ("N", C.c_int32), ("K", C.c_int32),
        ("lda", C.c_int32), ("ldb", C.c_int32), ("ldd", C.c_int32),
        ("d_fp32", C.c_int32), ("relu", C.c_int32), ("accumulate", C.c_int32), ("out_group_div", C.c_int32),
        ("group_inner", C.c_int32), ("a_gs2", C.c_int64), ("b_gs2", C.c_int64), ("d_gs2", C.c_int64),
        ("stats", C.c_void_p), ("stats_gs", C.c_int64),
    ]


class FbConv3x3(C.Structure):
    _fields_ = [
        ("x", C.c_void_p), ("w", C.c_void_p), ("y", C.c_void_p), ("w_gs", C.c_int64),
        ("G", C.c_int32), ("B", C.c_int32), ("H", C.c_int32), ("W", C.c_int32), ("cin", C.c_int32), ("cout", C.c_int32),
        ("stride", C.c_int32), ("dgrad", C.c_int32), ("stats", C.c_void_p), ("stats_gs", C.c_int64),
    ]


class FbConv3x3Wgrad(C.Structure):
    _fields_ = [
        ("x", C.c_void_p), ("dy", C.c_void_p), ("dw", C.c_void_p), ("dw_gs", C.c_int64),
        ("G", C.c_int32), ("B", C.c_int32), ("H", C.c_int32), ("W", C.c_int32), ("cin", C.c_int32), ("cout", C.c_int32),
        ("stride", C.c_int32), ("split", C.c_int32), ("accumulate", C.c_int32),
    ]


class FbDpStepParams(C.Structure):
    _fields_ = [
        ("peer_grad", C.c_void_p), ("peer_param", C.c_void_p), ("peer_flags", C.c_void_p), ("mom", C.c_void_p),
        ("n", C.c_int64), ("rank", C.c_int32), ("G", C.c_int32), ("epoch", C.c_uint32),
        ("lr", C.c_float), ("momentum", C.c_float), ("weight_decay", C.c_float), ("inv_world", C.c_float),
        ("nesterov", C.c_int32), ("max_blocks", C.c_int32), ("offset", C.c_int64),
    ]


class NativeError(RuntimeError):
    pass


def available() -> bool:
    return os.path.exists(LIB_PATH)


def lib():
    global _LIB
    if _LIB is not None:
        return _LIB
    with _LOCK:
        if _LIB is None:
            if not os.path.exists(LIB_PATH):
                raise NativeError(
                    f"{LIB_PATH} is missing: build it with `python -m fedml_b200.native.build` "
                    "(the GPU path has no PyTorch fallback)")
            L = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
            L.fb_last_error.restype = C.c_char_p
            for name in ("fb_round_flag_bytes", "fb_peer_barrier_flag_bytes", "fb_gram_workspace_bytes",
                         "fb_amp_normalize_workspace_bytes"):
                if hasattr(L, name):
                    getattr(L, name).restype = C.c_int64
            _LIB = L
    return _LIB


def check(rc: int, what: str = ""):
    if rc != 0:
        msg = lib().fb_last_error().decode("utf-8", "replace")
        raise NativeError(f"{what or 'native call'} failed (rc={rc}): {msg}")


def stream_ptr(device=None) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def ptr(t) -> int:
    if t is None:
        return 0
    return t.data_ptr()


# count of native kernel launches issued through this module (bench.py reports it as gpu_launches)
LAUNCHES = [0]


def count_launch(n: int = 1):
    LAUNCHES[0] += n
