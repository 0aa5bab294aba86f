"""B200-native (sm_100a) SAM segmentation branch of 2HandedAfforder: drop-in `visual_model`.

The directory name starts with a digit, so import it with
`importlib.import_module("2handedafforder_b200")`.
"""
from .build_sam import (_build_sam, build_sam, build_sam_vit_b, build_sam_vit_h, build_sam_vit_l,  # noqa: F401
                        sam_model_registry)
from .sam import B200Sam  # noqa: F401

__all__ = ["B200Sam", "build_sam", "build_sam_vit_h", "build_sam_vit_l", "build_sam_vit_b", "sam_model_registry",
           "_build_sam"]
