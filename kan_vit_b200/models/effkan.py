"""KANLinear (efficient-kan) -- drop-in for models/effkan.py:8-264 of the reference.

forward/backward run in libkanvit_b200 (B-spline basis evaluated in the GEMM prologue, SiLU
base branch in the same accumulator).  The maintenance methods that neither model.py nor
train.py of the reference ever call (b_splines, curve2coeff, update_grid,
regularization_loss) stay plain PyTorch, as SURVEY.md section 2 row 5 scopes them.
"""
import math

import torch

from .. import _lib as L
from ..ops import KanSpec, kan_apply


class KANLinear(torch.nn.Module):
    def __init__(self, in_features, out_features, grid_size=5, spline_order=3, scale_noise=0.1,
                 scale_base=1.0, scale_spline=1.0, enable_standalone_scale_spline=True,
                 base_activation=torch.nn.SiLU, grid_eps=0.02, grid_range=[-1, 1]):
        super().__init__()
        self.in_features = in_features
        self.out_features = out_features
        self.grid_size = grid_size
        self.spline_order = spline_order

        step = (grid_range[1] - grid_range[0]) / grid_size
        knots = torch.arange(-spline_order, grid_size + spline_order + 1) * step + grid_range[0]
        self.register_buffer("grid", knots.expand(in_features, -1).contiguous())

        self.base_weight = torch.nn.Parameter(torch.empty(out_features, in_features))
        self.spline_weight = torch.nn.Parameter(torch.empty(out_features, in_features, grid_size + spline_order))
        if enable_standalone_scale_spline:
            self.spline_scaler = torch.nn.Parameter(torch.empty(out_features, in_features))

        self.scale_noise = scale_noise
        self.scale_base = scale_base
        self.scale_spline = scale_spline
        self.enable_standalone_scale_spline = enable_standalone_scale_spline
        self.base_activation = base_activation()
        self.grid_eps = grid_eps
        self._spec = KanSpec(L.KV_EFFKAN, in_features, out_features, grid_size=grid_size,
                             spline_order=spline_order)
        self.reset_parameters()

    # ------------------------------------------------------------------ init / maintenance
    def reset_parameters(self):
        """Same draws, in the same order, as the reference (models/effkan.py:74-97)."""
        torch.nn.init.kaiming_uniform_(self.base_weight, a=math.sqrt(5) * self.scale_base)
        with torch.no_grad():
            noise = (torch.rand(self.grid_size + 1, self.in_features, self.out_features) - 0.5)
            noise = noise * self.scale_noise / self.grid_size
            pts = self.grid.T[self.spline_order:-self.spline_order]
            lead = 1.0 if self.enable_standalone_scale_spline else self.scale_spline
            self.spline_weight.data.copy_(lead * self.curve2coeff(pts, noise))
            if self.enable_standalone_scale_spline:
                torch.nn.init.kaiming_uniform_(self.spline_scaler, a=math.sqrt(5) * self.scale_spline)

    def b_splines(self, x: torch.Tensor):
        """(batch, in) -> (batch, in, grid_size + spline_order) Cox-de Boor bases, PyTorch."""
        assert x.dim() == 2 and x.size(1) == self.in_features
        t = self.grid
        xe = x.unsqueeze(-1)
        b = ((xe >= t[:, :-1]) & (xe < t[:, 1:])).to(x.dtype)
        for p in range(1, self.spline_order + 1):
            n = b.size(-1) - 1
            left = (xe - t[:, :n]) / (t[:, p:p + n] - t[:, :n])
            right = (t[:, p + 1:p + 1 + n] - xe) / (t[:, p + 1:p + 1 + n] - t[:, 1:1 + n])
            b = left * b[..., :n] + right * b[..., 1:1 + n]
        assert b.size() == (x.size(0), self.in_features, self.grid_size + self.spline_order)
        return b.contiguous()

    def curve2coeff(self, x: torch.Tensor, y: torch.Tensor):
        """Least-squares spline coefficients interpolating y (batch,in,out) at x (batch,in)."""
        assert x.dim() == 2 and x.size(1) == self.in_features
        assert y.size() == (x.size(0), self.in_features, self.out_features)
        design = self.b_splines(x).transpose(0, 1)          # (in, batch, coeff)
        target = y.transpose(0, 1)                          # (in, batch, out)
        coeff = torch.linalg.lstsq(design, target).solution  # (in, coeff, out)
        coeff = coeff.permute(2, 0, 1)
        assert coeff.size() == (self.out_features, self.in_features, self.grid_size + self.spline_order)
        return coeff.contiguous()

    @property
    def scaled_spline_weight(self):
        if self.enable_standalone_scale_spline:
            return self.spline_weight * self.spline_scaler.unsqueeze(-1)
        return self.spline_weight * 1.0

    @torch.no_grad()
    def update_grid(self, x: torch.Tensor, margin=0.01):
        """Re-fit the knots to the data distribution and refit the coefficients
        (behaviour of models/effkan.py:189-242)."""
        assert x.dim() == 2 and x.size(1) == self.in_features
        n = x.size(0)
        bases = self.b_splines(x).permute(1, 0, 2)                    # (in, batch, coeff)
        coeff = self.scaled_spline_weight.permute(1, 2, 0)            # (in, coeff, out)
        per_feature_out = torch.bmm(bases, coeff).permute(1, 0, 2)    # (batch, in, out)

        xs = torch.sort(x, dim=0)[0]
        pick = torch.linspace(0, n - 1, self.grid_size + 1, dtype=torch.int64, device=x.device)
        adaptive = xs[pick]
        step = (xs[-1] - xs[0] + 2 * margin) / self.grid_size
        idx = torch.arange(self.grid_size + 1, dtype=torch.float32, device=x.device).unsqueeze(1)
        uniform = idx * step + xs[0] - margin
        inner = self.grid_eps * uniform + (1 - self.grid_eps) * adaptive
        k = self.spline_order
        below = inner[:1] - step * torch.arange(k, 0, -1, device=x.device).unsqueeze(1)
        above = inner[-1:] + step * torch.arange(1, k + 1, device=x.device).unsqueeze(1)
        self.grid.copy_(torch.cat([below, inner, above], dim=0).T)
        self.spline_weight.data.copy_(self.curve2coeff(x, per_feature_out))

    def regularization_loss(self, regularize_activation=1.0, regularize_entropy=1.0):
        """L1 + entropy surrogate on the spline weights (models/effkan.py:244-264)."""
        l1 = self.spline_weight.abs().mean(-1)
        total = l1.sum()
        p = l1 / total
        entropy = -torch.sum(p * p.log())
        return regularize_activation * total + regularize_entropy * entropy

    # ------------------------------------------------------------------ hot path
    def grid_is_uniform(self) -> bool:
        """True when every feature still shares one uniform knot vector (the constructor's grid; it
        only changes through update_grid / load_state_dict).  The check reads the buffer once per
        version of ``grid`` (one device sync), so the kernels may use their closed-form prologue."""
        g = self.grid
        key = (g.data_ptr(), g._version, g.device)
        if getattr(self, "_uniform_key", None) != key:
            n = g.shape[1]
            t0, span = g[0, 0], g[0, -1] - g[0, 0]
            ideal = t0 + torch.arange(n, device=g.device, dtype=g.dtype) * (span / (n - 1))
            ok = bool(((g - ideal).abs().max() <= 1e-5 * span) & (span > 0))
            self._uniform_key, self._uniform = key, ok
            self._ends = (float(t0), float(g[0, -1]))
        return self._uniform

    def _uniform_flag(self) -> int:
        return L.KV_FLAG_UNIFORM_GRID if self.grid_is_uniform() else 0

    def grid_ends(self):
        """(first, last) knot of feature 0, cached with grid_is_uniform()."""
        self.grid_is_uniform()
        return self._ends

    def _slots(self):
        return [self.base_weight, self.spline_weight,
                self.spline_scaler if self.enable_standalone_scale_spline else None, self.grid]

    def forward(self, x: torch.Tensor):
        assert x.size(-1) == self.in_features
        if not isinstance(self.base_activation, torch.nn.SiLU):
            raise L.KanVitError("the fused kernel implements base_activation=SiLU only")
        shape = x.shape[:-1] + (self.out_features,)
        # a layer called on its own gets the closed-form uniform-grid prologue too (the model classes build their specs
        # through _dispatch.make_spec, which installs the same per-call flag); bound here so that copies of the module
        # look at their own grid
        self._spec.dyn_flags = self._uniform_flag
        y = kan_apply(x.reshape(-1, self.in_features), self._spec, [self._slots()])
        return y.reshape(shape)
