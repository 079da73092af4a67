"""kan_vit_b200 -- B200-native (sm_100a) implementation of the ViKANformer hot path behind the
reference's own nn.Module API (akshathmangudi/KAN-ViT: model.py, attention.py, models/*.py).

    from kan_vit_b200.model import VisionTransformer, TransformerBlock
    from kan_vit_b200.attention import MSA, FlashAttention
    from kan_vit_b200.models import KANLinear, ChebyKANLayer, FastKANLayer, SineKANLayer, NaiveFourierKANLayer
"""
from . import _lib  # noqa: F401
from .attention import MSA, FlashAttention
from .model import TransformerBlock, VisionTransformer
from .ops import get_precision, precision, set_precision
from .metrics import EpochMetrics
from .models import (ChebyKANLayer, FastKANLayer, KANLinear, NaiveFourierKANLayer, SineKANLayer)

__all__ = ["MSA", "FlashAttention", "TransformerBlock", "VisionTransformer", "KANLinear", "ChebyKANLayer", "FastKANLayer",
           "SineKANLayer", "NaiveFourierKANLayer", "set_precision", "get_precision", "precision", "EpochMetrics"]
