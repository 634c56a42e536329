"""Drop-in mirror of the reference's ``utils`` package for the hot path
(utils/utils.py, utils/trainClass.py::crop_image, utils/class_config.py::CFG)."""
from .utils import (custom_nms, Network_Wrapper, build_model, crop_image, classify_crops, box_dilate,  # noqa: F401
                    generate_annotation)
from .class_config import CFG  # noqa: F401
