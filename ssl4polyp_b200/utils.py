"""Model factories with the reference's names and arguments (/root/reference/src/ssl4polyp/utils/__init__.py:13-67)."""
from __future__ import annotations


def enable_deterministic_algorithms(*, warn_only: bool = True) -> None:
    import torch
    try:
        torch.use_deterministic_algorithms(True, warn_only=warn_only)
    except TypeError:
        torch.use_deterministic_algorithms(True)


def get_MAE_backbone(weight_path, head, num_classes, frozen, dense, out_token="cls"):
    """ViT-B/16 encoder of an MAE checkpoint + linear head (constants 768/12/12 as in the reference)."""
    from . import models
    return models.ViT_from_MAE(weight_path, head, num_classes, frozen, dense,
                               embed_dim=768, depth=12, num_heads=12, out_token=out_token)


def get_ImageNet_or_random_ViT(head, num_classes, frozen, dense, ImageNet_weights, out_token="cls"):
    from . import models
    return models.VisionTransformer_from_Any(head, num_classes, frozen, dense, 768, 12, 12, out_token, ImageNet_weights)
