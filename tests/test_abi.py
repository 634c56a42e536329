"""CPU tests of the drop-in boundary: the C-ABI library builds for sm_100a, loads, and exports
every symbol include/yv8vit.h declares; the host mirror keeps the reference's names and refuses
to run without a GPU (no CPU fallback)."""
import ctypes
import os
import re
import subprocess

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib_path():
    from yolov8_vit_b200 import build
    return build.build()


def test_header_symbols_exported(lib_path):
    hdr = open(os.path.join(ROOT, "include", "yv8vit.h")).read()
    declared = set(re.findall(r"\b(yv_[a-z0-9_]+)\s*\(", hdr))
    lib = ctypes.CDLL(lib_path)
    for sym in sorted(declared):
        assert hasattr(lib, sym), f"{sym} declared in yv8vit.h but not exported"
    from yolov8_vit_b200 import _lib
    assert declared == set(_lib.SYMBOLS)


def test_library_is_sm100a_with_tcgen05(lib_path):
    sass = subprocess.run(["cuobjdump", "-sass", lib_path], capture_output=True, text=True).stdout
    assert "sm_100a" in sass
    for mnemonic in ("UTCHMMA", "UTMALDG", "LDTM"):      # tcgen05.mma, TMA load, tcgen05.ld
        assert mnemonic in sass, mnemonic


def test_no_gpu_calls_without_device(lib_path):
    from yolov8_vit_b200 import _lib
    lib = _lib.load()
    assert lib.yv_version() == 100
    assert lib.yv_decode_filter_workspace_bytes(64, 80, 8400) > 64 * 8400 * 28
    assert lib.yv_nms_workspace_bytes(2, 1000) >= 2 * 1000 * 56
    assert isinstance(lib.yv_last_error(), bytes)


def test_reference_call_surface_and_no_cpu_fallback(lib_path):
    import inspect
    import yolov8_vit_b200 as pkg
    from yolov8_vit_b200.utils import utils as U
    assert list(inspect.signature(U.custom_nms).parameters)[:3] == ["boxes", "scores", "iou_threshold"]
    assert inspect.signature(U.custom_nms).parameters["iou_threshold"].default == 0.45
    assert list(inspect.signature(U.crop_image).parameters)[:6] == ["image_path", "x_min", "y_min", "x_max", "y_max", "training"]
    assert list(inspect.signature(U.build_model).parameters)[:3] == ["CFG", "modelName", "pretrained_path"]
    assert "pretrained" in inspect.signature(U.build_model).parameters      # app.py:35 spelling
    assert hasattr(U.Network_Wrapper, "forward")
    if not torch.cuda.is_available():
        from yolov8_vit_b200 import ops
        with pytest.raises(Exception):
            ops.nms_batched(torch.zeros(1, 4, 4), torch.zeros(1, 4))
        with pytest.raises(Exception):
            ops.decode_filter(torch.zeros(1, 84, 8400))


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "yolov8_vit_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh")):
                src = open(os.path.join(dp, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", src, re.M), f
