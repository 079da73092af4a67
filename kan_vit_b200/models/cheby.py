"""ChebyKANLayer -- drop-in for models/cheby.py:10-48 of the reference."""
import torch
import torch.nn as nn

from .. import _lib as L
from ..ops import KanSpec, kan_apply


class ChebyKANLayer(nn.Module):
    """Y[t,o] = sum_i sum_n T_n(tanh x[t,i]) * cheby_coeffs[i,o,n]; output is always 2-D
    (reference models/cheby.py:38,47)."""

    def __init__(self, input_dim, output_dim, degree):
        super().__init__()
        self.inputdim = input_dim
        self.outdim = output_dim
        self.degree = degree
        coeffs = torch.empty(input_dim, output_dim, degree + 1)
        nn.init.normal_(coeffs, mean=0.0, std=1 / (input_dim * (degree + 1)))
        self.cheby_coeffs = nn.Parameter(coeffs)
        self.register_buffer("arange", torch.arange(0, degree + 1, 1))
        self._spec = KanSpec(L.KV_CHEBY, input_dim, output_dim, grid_size=degree)

    def _slots(self):
        return [self.cheby_coeffs]

    def forward(self, x):
        x2 = x.reshape(-1, self.inputdim)
        return kan_apply(x2, self._spec, [self._slots()])
