"""yolov8_vit_b200 -- B200-native (sm_100a) second-stage hot path of Voyager0587/yolov8-vit.

head [B,4+nc,A] -> conf filter -> custom_nms -> dilation -> crop/resize/normalise -> ViT-B/16.
All compute is in libyv8vit.so (hand-written CUDA, C-ABI in include/yv8vit.h); this package is the
host-side mirror of the reference's Python call surface.  No CPU fallback.
"""
from .pipeline import HotPath, HotPathConfig, HotPathResult, PendingHost, records_to_tuples  # noqa: F401
from .vit import ViTConfig, ViTModel  # noqa: F401
from .utils.utils import custom_nms, Network_Wrapper, build_model, crop_image  # noqa: F401

__all__ = ["HotPath", "HotPathConfig", "HotPathResult", "records_to_tuples", "ViTConfig", "ViTModel",
           "custom_nms", "Network_Wrapper", "build_model", "crop_image"]
