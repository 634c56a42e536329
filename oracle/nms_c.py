"""ctypes loader for oracle/nms_c.c (TEST INFRASTRUCTURE ONLY, see oracle/__init__.py)."""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liboracle_nms.so")
_lib = None


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "nms_c.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["gcc", "-O2", "-ffp-contract=off", "-shared", "-fPIC", src, "-o", _SO])
    return _SO


def lib():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(build())
        _lib.yv_oracle_nms.restype = ctypes.c_int
        _lib.yv_oracle_nms.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int,
                                       ctypes.c_float, ctypes.c_void_p]
    return _lib


def nms(boxes, scores, iou_threshold: float = 0.45):
    """boxes [N,4] f32, scores [N] f32 (numpy or torch-CPU) -> list[int]."""
    b = np.ascontiguousarray(np.asarray(boxes, dtype=np.float32))
    s = np.ascontiguousarray(np.asarray(scores, dtype=np.float32))
    n = int(s.shape[0])
    keep = np.empty(max(n, 1), dtype=np.int32)
    k = lib().yv_oracle_nms(b.ctypes.data, s.ctypes.data, n,
                            ctypes.c_float(np.float32(iou_threshold)), keep.ctypes.data)
    return keep[:k].tolist()
