"""FastKANLayer -- drop-in for models/fastkan.py:6-76 of the reference."""
import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import _lib as L
from ..ops import KanSpec, kan_apply


class SplineLinear(nn.Linear):
    """Bias-free Linear with truncated-normal init (models/fastkan.py:6-12)."""

    def __init__(self, in_features: int, out_features: int, init_scale: float = 0.1, **kw) -> None:
        self.init_scale = init_scale
        super().__init__(in_features, out_features, bias=False, **kw)

    def reset_parameters(self) -> None:
        nn.init.trunc_normal_(self.weight, mean=0, std=self.init_scale)


class RadialBasisFunction(nn.Module):
    """Holder of the RBF centres ``grid`` (a no-grad Parameter, as in the reference) and the
    width ``denominator`` (models/fastkan.py:15-30).  Standalone calls evaluate in PyTorch; inside
    FastKANLayer the RBFs are evaluated in the GEMM prologue."""

    def __init__(self, grid_min: float = -2., grid_max: float = 2., num_grids: int = 8,
                 denominator: float = None):
        super().__init__()
        self.grid = torch.nn.Parameter(torch.linspace(grid_min, grid_max, num_grids), requires_grad=False)
        self.denominator = denominator or (grid_max - grid_min) / (num_grids - 1)

    def forward(self, x):
        return torch.exp(-((x[..., None] - self.grid) / self.denominator) ** 2)


class FastKANLayer(nn.Module):
    """Y = W_s vec(exp(-((LN(x)_i - g_j)/h)^2)) + W_b silu(x) + b_b."""

    def __init__(self, input_dim: int, output_dim: int, grid_min: float = -2., grid_max: float = 2.,
                 num_grids: int = 8, use_base_update: bool = True, base_activation=F.silu,
                 spline_weight_init_scale: float = 0.1) -> None:
        super().__init__()
        self.input_dim, self.output_dim, self.num_grids = input_dim, output_dim, num_grids
        self.layernorm = nn.LayerNorm(input_dim)
        self.rbf = RadialBasisFunction(grid_min, grid_max, num_grids)
        self.spline_linear = SplineLinear(input_dim * num_grids, output_dim, spline_weight_init_scale)
        self.use_base_update = use_base_update
        if use_base_update:
            self.base_activation = base_activation
            self.base_linear = nn.Linear(input_dim, output_dim)
        self._specs = {}

    def _spec(self, time_benchmark):
        fused_base = self.use_base_update and self.base_activation is F.silu
        flags = (0 if fused_base else L.KV_FLAG_NO_BASE) | (L.KV_FLAG_NO_LAYERNORM if time_benchmark else 0)
        key = (flags, float(self.rbf.denominator))
        if key not in self._specs:
            self._specs[key] = KanSpec(L.KV_FASTKAN, self.input_dim, self.output_dim,
                                       grid_size=self.num_grids, flags=flags,
                                       rbf_denom=self.rbf.denominator)
        return self._specs[key], fused_base

    def _slots(self, fused_base=True):
        base_w = self.base_linear.weight if fused_base else None
        base_b = self.base_linear.bias if fused_base else None
        return [self.layernorm.weight, self.layernorm.bias, self.rbf.grid, self.spline_linear.weight,
                base_w, base_b]

    def forward(self, x, time_benchmark=False):
        spec, fused_base = self._spec(time_benchmark)
        shape = x.shape[:-1] + (self.output_dim,)
        y = kan_apply(x.reshape(-1, self.input_dim), spec, [self._slots(fused_base)]).reshape(shape)
        if self.use_base_update and not fused_base:  # non-SiLU activation: evaluated outside the kernel
            y = y + self.base_linear(self.base_activation(x))
        return y
