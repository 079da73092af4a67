"""NaiveFourierKANLayer -- drop-in for models/nfkan.py:5-52 of the reference."""
import numpy
import torch

from .. import _lib as L
from ..ops import KanSpec, kan_apply


class NaiveFourierKANLayer(torch.nn.Module):
    """Y = sum_i sum_{k=1..G} cos(k x_i) F[0,o,i,k] + sin(k x_i) F[1,o,i,k] (+ bias).
    The kwarg spelling ``smootorch_initialization`` is the reference's (models/nfkan.py:6)."""

    def __init__(self, inputdim, outdim, gridsize, addbias=True, smootorch_initialization=False):
        super().__init__()
        self.gridsize = gridsize
        self.addbias = addbias
        self.inputdim = inputdim
        self.outdim = outdim
        if smootorch_initialization:
            norm = (torch.arange(gridsize) + 1) ** 2
        else:
            norm = numpy.sqrt(gridsize)
        self.fouriercoeffs = torch.nn.Parameter(
            torch.randn(2, outdim, inputdim, gridsize) / (numpy.sqrt(inputdim) * norm))
        if addbias:
            self.bias = torch.nn.Parameter(torch.zeros(1, outdim))
        self._spec = KanSpec(L.KV_FOURIER, inputdim, outdim, grid_size=gridsize)

    def _slots(self):
        return [self.fouriercoeffs, self.bias if self.addbias else None]

    def forward(self, x):
        shape = x.shape[:-1] + (self.outdim,)
        y = kan_apply(x.reshape(-1, self.inputdim), self._spec, [self._slots()])
        return y.reshape(shape)
