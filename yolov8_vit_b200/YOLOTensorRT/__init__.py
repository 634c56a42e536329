"""Mirror of the reference's ``YOLOTensorRT`` detector wrapper (an un-vendored submodule upstream,
described in YOLOTensorRT_yolodet_py_解读.md): ``inferdet.main`` / ``det_postprocess`` with the same
signatures, running the batched CUDA hot path.  The TensorRT engine itself is out of scope: ``main``
takes any callable ``Engine(blob) -> head`` (see ``models.HeadEngine``)."""
from .inferdet import main, det_postprocess, draw_image, CLASSES, COLORS  # noqa: F401
