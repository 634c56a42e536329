"""CPU oracle for the yolov8-vit second-stage hot path.

TEST INFRASTRUCTURE ONLY. Nothing under ``oracle/`` is part of the product:
only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline``
/ ``--impl reference`` legs may import it, and only as the checker / the timed
CPU baseline.  The product path (``yolov8-vit_b200``) never imports this
package and fails loudly when its CUDA library is missing.

What pins it: the reference (Voyager0587/yolov8-vit) ships no tests, golden vectors or
fixtures for this path, ``custom_nms`` exists only as a markdown
snippet, the detector wrapper is an un-vendored submodule and ``timm`` is absent -- but
the reference's OWN TEXT for ``custom_nms`` (README.md:62-83), ``crop_image``
(utils/trainClass.py:70-93), ``Network_Wrapper`` (utils/utils.py:59-72) and the
Detect-head decode (docs/YOLO_TensorRT_Technical.md:14-29,72-76) is extracted byte for
byte and executed by ``oracle/make_ref.py`` (-> git-ignored ``oracle/_ref/``), and
``tests/test_ref.py`` holds the restatements below to it.  Stages the reference leaves to
libraries it does not pin are checked against those libraries (cv2 nearest resize,
torch bilinear interpolate, torchvision ViT).  Rows without either ("oracle only": the
a1 conf filter + cxcywh decode, whose only upstream form is a C++ excerpt) are listed in
DESIGN.md §2.  The golden vectors under ``tests/golden/`` are generated from the oracle by
``tests/golden/make_golden.py`` with fixed seeds.
"""
from .hotpath import (  # noqa: F401
    custom_nms, decode_head, round_boxes, dilate_box, crop_resize_norm,
    nearest_index_table, patchify, ViTConfig, make_vit_weights, vit_forward,
    pipeline_image,
)
