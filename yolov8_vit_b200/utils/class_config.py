"""Config surface read by build_model (mirrors utils/class_config.py:12-24 of the reference)."""


class CFG:
    device = "cuda:0"
    img_size = 224
    num_classes = 5
    # north star pins ViT-B/16; the reference-literal name is
    # 'vit_base_patch8_224.augreg_in21k' (utils/class_config.py:21), accepted by build_model too.
    modelName = "vit_base_patch16_224.augreg_in21k"
    classes = ["good", "broke", "lose", "uncovered", "circle"]
