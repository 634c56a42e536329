"""CPU oracle for the yolov8-vit second-stage hot path.

TEST INFRASTRUCTURE ONLY. Nothing under ``oracle/`` is part of the product:
only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline``
/ ``--impl reference`` legs may import it, and only as the checker / the timed
CPU baseline.  The product path (``yolov8-vit_b200``) never imports this
package and fails loudly when its CUDA library is missing.

PARITY UNPINNED: the reference (Voyager0587/yolov8-vit) ships no tests, golden
vectors or fixtures for this path; ``custom_nms`` exists
only as a markdown snippet, the detector wrapper is an un-vendored submodule
and ``timm`` is absent.  The oracle therefore restates the snippets verbatim
(each function cites the reference file:line it follows) and the golden
vectors under ``tests/golden/`` are generated from it by
``tests/golden/make_golden.py`` with fixed seeds.
"""
from .hotpath import (  # noqa: F401
    custom_nms, decode_head, round_boxes, dilate_box, crop_resize_norm,
    nearest_index_table, patchify, ViTConfig, make_vit_weights, vit_forward,
    pipeline_image,
)
