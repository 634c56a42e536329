"""``YOLOTensorRT.models`` surface used by the reference's call sites (app.py:17,28-29):
``TRTModule(engine, device)`` + ``set_desired([...])``.  The TensorRT runtime is not part of the hot path
(and is not installed here); ``HeadEngine`` is the protocol ``inferdet.main`` needs from a detector."""
from __future__ import annotations

import torch


class HeadEngine:
    """Wraps any callable ``f(blob [B,3,H,W] f32) -> head`` so it looks like the reference's ``TRTModule``:
    ``inp_info[0].shape`` gives the input size, ``set_desired`` is accepted.  ``kind`` says what the callable
    returns: 'head' = decoded ``[B,4+nc,A]`` (cx,cy,w,h + probabilities, row a1), 'raw' = Detect-head raw
    ``[B,64+nc,A]`` (row f1), 'nms' = the engine's EfficientNMS tuple (num_dets, bboxes, scores, labels)."""

    class _Info:
        def __init__(self, shape):
            self.shape = shape

    def __init__(self, fn, input_hw=(640, 640), kind: str = "head", levels=None):
        assert kind in ("head", "raw", "nms")
        self.fn, self.kind, self.levels = fn, kind, levels
        self.inp_info = [HeadEngine._Info((1, 3, int(input_hw[0]), int(input_hw[1])))]
        self.desired = None

    def set_desired(self, names):
        self.desired = list(names)

    def __call__(self, blob: torch.Tensor):
        return self.fn(blob)


class TRTModule(HeadEngine):
    """``TRTModule(engine_path, device)`` (app.py:28).  Deserialising a TensorRT engine needs the
    ``tensorrt`` package, which this repo neither ships nor replaces: constructing it without one raises."""

    def __init__(self, weight, device=None):
        try:
            import tensorrt  # noqa: F401
        except ImportError as e:
            raise RuntimeError("TRTModule needs the tensorrt runtime (out of scope of the hot path); wrap your "
                               "detector in yolov8_vit_b200.YOLOTensorRT.models.HeadEngine instead") from e
        raise NotImplementedError("TensorRT engine execution is outside SURVEY.md §8's hot path")
