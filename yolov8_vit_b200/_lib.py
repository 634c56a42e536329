"""ctypes binding of libyv8vit.so (the C-ABI declared in include/yv8vit.h).

PyTorch is used only as plumbing: device buffers, streams, torch.distributed.
There is NO CPU fallback: if the library is missing or no sm_100 GPU is
present, the ops raise.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_char_p, c_float, c_int, c_size_t, c_void_p, POINTER

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libyv8vit.so")

# every symbol include/yv8vit.h declares (tests check the .so exports all of them)
SYMBOLS = [
    "yv_version", "yv_last_error", "yv_check_device",
    "yv_decode_filter_workspace_bytes", "yv_decode_filter", "yv_decode_filter_raw", "yv_letterbox_geometry", "yv_letterbox_blob",
    "yv_nms_workspace_bytes", "yv_nms_batched", "yv_nms_batched_rule", "yv_nms_batched_noalloc",
    "yv_dilate_boxes", "yv_build_crops", "yv_crop_resize",
    "yv_vit_create", "yv_vit_destroy", "yv_vit_set_param", "yv_vit_missing_params", "yv_vit_last_block_cls_rows",
    "yv_vit_workspace_bytes", "yv_vit_forward",
    "yv_tune_sm_budget", "yv_gemm_bf16", "yv_linear_ln_f16_ws_bytes", "yv_linear_ln_f16", "yv_residual_linear_f16_ws_bytes", "yv_residual_linear_f16", "yv_layernorm_bf16", "yv_attention_bf16", "yv_pack_results",
    "yv_prof_enable", "yv_prof_kinds", "yv_prof_kind_name", "yv_prof_read", "yv_debug_attn_trace",
]


class YvError(RuntimeError):
    pass


class VitConfigC(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int32) for n in
                ("img", "patch", "dim", "depth", "heads", "mlp", "head_out", "hidden", "num_classes")] + \
               [("eps", c_float)]


_lib = None


def load() -> ctypes.CDLL:
    """Loads the library (never builds it: run ``python -m yolov8_vit_b200.build``)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise YvError(f"{LIB_PATH} is missing: build it with `python -m yolov8_vit_b200.build` "
                      "(there is no CPU fallback for this path)")
    lib = ctypes.CDLL(LIB_PATH)
    lib.yv_version.restype = c_int
    lib.yv_last_error.restype = c_char_p
    lib.yv_check_device.argtypes = [c_int]
    lib.yv_decode_filter_workspace_bytes.restype = c_size_t
    lib.yv_decode_filter_workspace_bytes.argtypes = [c_int, c_int, c_int]
    lib.yv_decode_filter.argtypes = [c_void_p, c_int, c_int, c_int, c_float, c_int, c_float, c_float, c_float,
                                     c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t,
                                     c_void_p]
    lib.yv_decode_filter_raw.argtypes = [c_void_p, c_int, c_int, c_void_p, c_void_p, c_float, c_int, c_float, c_float,
                                         c_float, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                         c_size_t, c_void_p]
    lib.yv_letterbox_geometry.argtypes = [c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]
    lib.yv_letterbox_blob.argtypes = [c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p]
    lib.yv_nms_workspace_bytes.restype = c_size_t
    lib.yv_nms_workspace_bytes.argtypes = [c_int, c_int]
    lib.yv_nms_batched.argtypes = [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_float, c_int, c_void_p,
                                   c_void_p, c_void_p, c_size_t, c_void_p]
    lib.yv_nms_batched_rule.argtypes = [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_float, c_int, c_int,
                                        c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]
    lib.yv_nms_batched_noalloc.argtypes = [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_float, c_int, c_void_p,
                                           c_void_p, c_void_p]
    lib.yv_dilate_boxes.argtypes = [c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p]
    lib.yv_build_crops.argtypes = [c_void_p] * 6 + [c_int] * 8 + [c_void_p] * 10 + [c_int, c_void_p, c_void_p]
    lib.yv_crop_resize.argtypes = [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_int, c_void_p, c_int,
                                   c_int, c_int, c_int, c_void_p, c_void_p]
    lib.yv_vit_create.argtypes = [POINTER(VitConfigC), POINTER(c_void_p)]
    lib.yv_vit_destroy.argtypes = [c_void_p]
    lib.yv_vit_destroy.restype = None
    lib.yv_vit_set_param.argtypes = [c_void_p, c_char_p, c_void_p, c_size_t, c_void_p]
    lib.yv_vit_missing_params.argtypes = [c_void_p]
    lib.yv_vit_last_block_cls_rows.argtypes = [c_void_p, c_int]
    lib.yv_vit_workspace_bytes.restype = c_size_t
    lib.yv_vit_workspace_bytes.argtypes = [c_void_p, c_int]
    lib.yv_vit_forward.argtypes = [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t,
                                   c_void_p]
    lib.yv_gemm_bf16.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int,
                                 c_int, c_void_p]
    lib.yv_linear_ln_f16_ws_bytes.argtypes = [c_int, c_int, c_int]
    lib.yv_linear_ln_f16_ws_bytes.restype = ctypes.c_size_t
    lib.yv_linear_ln_f16.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int,
                                     c_float, c_int, c_void_p, ctypes.c_size_t, c_void_p]
    lib.yv_tune_sm_budget.argtypes = [c_int, c_int]
    lib.yv_residual_linear_f16_ws_bytes.argtypes = [c_int, c_int]
    lib.yv_residual_linear_f16_ws_bytes.restype = c_size_t
    lib.yv_residual_linear_f16.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_float,
                                           c_void_p, c_void_p, c_size_t, c_void_p]
    lib.yv_layernorm_bf16.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_float, c_void_p]
    lib.yv_attention_bf16.argtypes = [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p]
    lib.yv_pack_results.argtypes = [c_void_p] * 6 + [c_int, c_int, c_void_p, c_void_p]
    lib.yv_prof_enable.argtypes = [c_int]
    lib.yv_prof_kind_name.restype = c_char_p
    lib.yv_prof_kind_name.argtypes = [c_int]
    lib.yv_prof_read.argtypes = [c_void_p, c_void_p, c_void_p]
    _lib = lib
    return lib


def prof_enable(on: bool) -> None:
    check(load().yv_prof_enable(1 if on else 0), "yv_prof_enable")


def prof_read() -> dict:
    """{kernel class: {"ms": device ms (event-bracketed), "launches": n, "work": bytes or flops}}"""
    lib = load()
    k = lib.yv_prof_kinds()
    ms = (ctypes.c_double * k)()
    ln = (ctypes.c_longlong * k)()
    wk = (ctypes.c_double * k)()
    check(lib.yv_prof_read(ms, ln, wk), "yv_prof_read")
    return {lib.yv_prof_kind_name(i).decode(): {"ms": ms[i], "launches": int(ln[i]), "work": wk[i]} for i in range(k)}


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        raise YvError(f"{what or 'libyv8vit'} failed ({rc}): {load().yv_last_error().decode()}")


_checked_devices = set()


def require_device(t: torch.Tensor) -> int:
    """The tensor must live on an sm_100 CUDA device; returns the device index."""
    if not t.is_cuda:
        raise YvError("yolov8_vit_b200 runs on a B200 (sm_100a) only: got a CPU tensor and there is no CPU fallback")
    dev = t.device.index if t.device.index is not None else torch.cuda.current_device()
    if dev not in _checked_devices:
        check(load().yv_check_device(dev), "yv_check_device")
        _checked_devices.add(dev)
    return dev


def ptr(t) -> c_void_p:
    return c_void_p(0 if t is None else t.data_ptr())


def stream_ptr(device=None) -> c_void_p:
    return c_void_p(torch.cuda.current_stream(device).cuda_stream)
