"""ctypes binding of libbsc_hot.so (include/bsc_hot.h).

There is no CPU fallback: if the shared object is missing or the device is not sm_100 the import of
the ops fails with an exception that says so.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, byref, c_char_p, c_float, c_int, c_int32, c_int64, c_size_t, c_void_p

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("BSC_LIB_PATH") or os.path.join(_HERE, "libbsc_hot.so")  # override: A/B of two builds

BSC_F32, BSC_BF16, BSC_F16, BSC_F64 = 0, 1, 2, 3
MODE_F32, MODE_TC16, MODE_TC16X3 = 0, 1, 2
TC_MODES = (MODE_TC16, MODE_TC16X3)
BWD_AUTO, BWD_RECOMPUTE, BWD_BLOCKS, BWD_SCATTER = -1, 0, 3, 4
BWD_NAMES = {"auto": BWD_AUTO, "recompute": BWD_RECOMPUTE, "blocks": BWD_BLOCKS, "scatter": BWD_SCATTER}
ABI_VERSION = 200
NORM_TORCH, NORM_SKLEARN = 0, 1

EXPORTS = [
    "bsc_version", "bsc_last_error", "bsc_device_check", "bsc_l2norm_fwd", "bsc_rows_to_f16", "bsc_rows_to_f16x3", "bsc_label_stats", "bsc_label_stats_workspace_bytes",
    "bsc_label_stats_static",
    "bsc_gather_layout", "bsc_gather_push", "bsc_gather_push_labels", "bsc_gather_ring",
    "bsc_infonce_stats_len", "bsc_infonce_aux_len", "bsc_infonce_loss_scratch_bytes", "bsc_infonce_plan", "bsc_infonce_fwd", "bsc_infonce_loss",
    "bsc_infonce_bwd_partials", "bsc_infonce_bwd_finish", "bsc_bank_search_workspace_bytes", "bsc_bank_search", "bsc_bank_screen_parts", "bsc_bank_screen", "bsc_bank_rescore", "bsc_topk_merge", "bsc_taxonomy_eval",
    "bsc_set_timing", "bsc_last_kernel_ms", "bsc_launch_count", "bsc_last_bwd_mode",
]


class BscError(RuntimeError):
    pass


class InfoNceDesc(ctypes.Structure):
    _fields_ = [
        ("n_global", c_int64), ("n_local", c_int64), ("row0", c_int64),
        ("d", c_int32), ("mode", c_int32), ("logit_scale", c_float), ("x16_dtype", c_int32),
        ("xh", c_void_p * 3), ("inv", c_void_p * 3), ("r", c_void_p), ("rep", c_void_p), ("next", c_void_p), ("rmax", c_void_p),
        ("bwd_mode", c_int32), ("world", c_int32), ("rank", c_int32), ("reserved", c_int32),
    ]


class InfoNceExchange(ctypes.Structure):
    _fields_ = [("stats_peers", POINTER(c_void_p)), ("stats_multicast", c_void_p), ("arrive", c_void_p),
                ("arrive_target", ctypes.c_uint64)]


class InfoNcePlan(ctypes.Structure):
    _fields_ = [("bwd_mode", c_int32), ("reserved", c_int32), ("fwd_workspace_bytes", c_size_t),
                ("bwd_workspace_bytes", c_size_t), ("exchange_bytes", c_size_t)]


_lib = None


def lib():
    """Load the library once; raise if it was not built (run ``python __graft_entry__.py`` / build.py)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise BscError(
            f"{LIB_PATH} is missing: build it with `python bioscan-clip_b200/build.py` (needs nvcc, sm_100a). "
            "There is no CPU fallback for the bsc ops.")
    L = ctypes.CDLL(LIB_PATH)
    L.bsc_version.restype = c_int
    L.bsc_last_error.restype = c_char_p
    L.bsc_device_check.restype = c_int
    L.bsc_l2norm_fwd.argtypes = [c_void_p, c_int, c_int64, c_int, c_int64, c_void_p, c_int, c_void_p, c_void_p,
                                 c_int, c_void_p]
    L.bsc_rows_to_f16.argtypes = [c_void_p, c_int, c_int64, c_int, c_int64, c_void_p, c_void_p, c_void_p]
    L.bsc_rows_to_f16x3.argtypes = [c_void_p, c_int, c_int64, c_int, c_int64, c_void_p, c_void_p, c_void_p]
    L.bsc_label_stats.argtypes = [c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t,
                                  c_void_p]
    L.bsc_label_stats_workspace_bytes.argtypes = [c_int64, POINTER(c_size_t)]
    L.bsc_label_stats_static.argtypes = [c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]
    L.bsc_bank_screen_parts.argtypes = [c_int64, c_int64, c_int, POINTER(c_int)]
    L.bsc_bank_screen.argtypes = [c_void_p, c_int64, c_void_p, c_int64, c_int, c_int64, c_int, c_void_p, c_void_p, c_void_p]
    L.bsc_bank_rescore.argtypes = [c_void_p, c_int64, c_void_p, c_int64, c_int, c_int64, c_void_p, c_void_p, c_int, c_int,
                                   c_float, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]
    L.bsc_taxonomy_eval.argtypes = [c_void_p, c_int64, c_int, c_void_p, c_int64, c_void_p, c_int, POINTER(c_int32), c_int,
                                    POINTER(c_int64), c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]
    L.bsc_gather_layout.argtypes = [c_int64, c_int, c_int, POINTER(c_size_t), POINTER(c_size_t), POINTER(c_size_t),
                                    POINTER(c_size_t), POINTER(c_size_t), POINTER(c_size_t)]
    L.bsc_gather_push_labels.argtypes = [c_void_p, c_int64, c_int64, c_int64, c_int, POINTER(c_void_p), c_void_p, c_int,
                                         c_void_p]
    L.bsc_gather_push.argtypes = [POINTER(c_void_p), c_int, POINTER(c_int64), c_void_p, c_int64, c_int64, c_int64, c_int,
                                  POINTER(c_void_p), c_void_p, c_int, c_int, c_int, ctypes.c_uint64, c_void_p]
    L.bsc_gather_ring.argtypes = [POINTER(c_void_p), c_int, c_int, c_int64, c_int64, c_int64, c_int, ctypes.c_uint,
                                  c_void_p]
    L.bsc_infonce_aux_len.argtypes = [c_int64]
    L.bsc_infonce_aux_len.restype = c_int64
    L.bsc_infonce_stats_len.argtypes = [c_int64]
    L.bsc_infonce_stats_len.restype = c_int64
    L.bsc_infonce_plan.argtypes = [POINTER(InfoNceDesc), c_int, c_int, c_size_t, POINTER(InfoNcePlan)]
    L.bsc_infonce_fwd.argtypes = [POINTER(InfoNceDesc), c_void_p, POINTER(InfoNceExchange), c_void_p, c_size_t, c_void_p]
    L.bsc_infonce_loss.argtypes = [POINTER(InfoNceDesc), c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]
    L.bsc_infonce_loss_scratch_bytes.argtypes = []
    L.bsc_infonce_loss_scratch_bytes.restype = c_size_t
    L.bsc_infonce_bwd_partials.argtypes = [POINTER(InfoNceDesc), c_void_p, POINTER(c_void_p), c_void_p, c_size_t,
                                           c_void_p]
    L.bsc_infonce_bwd_finish.argtypes = [POINTER(InfoNceDesc), c_void_p, POINTER(c_void_p), c_int, POINTER(c_int64),
                                         c_void_p, POINTER(c_void_p), c_void_p, c_void_p, c_size_t, c_void_p]
    L.bsc_bank_search_workspace_bytes.argtypes = [c_int64, c_int64, c_int, c_int, c_int, POINTER(c_size_t)]
    L.bsc_bank_search.argtypes = [c_void_p, c_int64, c_void_p, c_int64, c_int, c_int, c_int64, c_int, c_void_p,
                                  c_void_p, c_void_p, c_size_t, c_void_p]
    L.bsc_topk_merge.argtypes = [c_void_p, c_void_p, c_int, c_int64, c_int, c_void_p, c_void_p, c_void_p]
    L.bsc_set_timing.argtypes = [c_int]
    L.bsc_last_kernel_ms.argtypes = [c_int]
    L.bsc_last_kernel_ms.restype = c_float
    L.bsc_last_bwd_mode.argtypes = []
    L.bsc_last_bwd_mode.restype = c_int
    L.bsc_launch_count.argtypes = [c_int]
    L.bsc_launch_count.restype = c_int64
    for name in EXPORTS:
        getattr(L, name)  # AttributeError here means header and library disagree
    _lib = L
    return L


def check(rc: int) -> None:
    if rc != 0:
        msg = lib().bsc_last_error().decode("utf-8", "replace")
        raise BscError(f"libbsc_hot error {rc}: {msg}")


_checked_devices = set()


def require_device(device: torch.device) -> None:
    """The ops run on sm_100 CUDA devices only; anything else is an error, never a fallback."""
    if device.type != "cuda":
        raise BscError(f"bsc ops need CUDA tensors on a B200 (got a tensor on '{device}'); there is no CPU path")
    idx = device.index if device.index is not None else torch.cuda.current_device()
    if idx in _checked_devices:
        return
    with torch.cuda.device(idx):
        check(lib().bsc_device_check())
    _checked_devices.add(idx)


def dtype_code(dt: torch.dtype) -> int:
    if dt == torch.float32:
        return BSC_F32
    if dt == torch.bfloat16:
        return BSC_BF16
    if dt == torch.float16:
        return BSC_F16
    if dt == torch.float64:
        return BSC_F64  # accepted by bsc_l2norm_fwd only (the reference's float64 eval features)
    raise BscError(f"unsupported dtype {dt}: bsc ops take float32, bfloat16 or float16")


def ptr(t) -> int:
    return 0 if t is None else t.data_ptr()


def stream_ptr(device) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def launch_count(reset: bool = False) -> int:
    return int(lib().bsc_launch_count(1 if reset else 0))


def set_timing(on: bool) -> None:
    lib().bsc_set_timing(1 if on else 0)


def last_bwd_mode() -> int:
    return int(lib().bsc_last_bwd_mode())


def last_kernel_ms(which: int) -> float:
    return float(lib().bsc_last_kernel_ms(which))
