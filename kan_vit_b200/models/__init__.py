"""KAN layer modules with the reference's class names, constructor arguments and parameter
names (models/*.py of akshathmangudi/KAN-ViT); forward/backward run in libkanvit_b200."""
from .cheby import ChebyKANLayer
from .effkan import KANLinear
from .fastkan import FastKANLayer, RadialBasisFunction, SplineLinear
from .nfkan import NaiveFourierKANLayer
from .sinekan import SineKANLayer, forward_step

__all__ = ["KANLinear", "ChebyKANLayer", "FastKANLayer", "RadialBasisFunction", "SplineLinear",
           "SineKANLayer", "forward_step", "NaiveFourierKANLayer"]
