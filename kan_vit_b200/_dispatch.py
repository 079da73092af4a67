"""Maps the reference-named modules onto kernel specs (family, sizes, pointer-table slots)."""
import torch.nn as nn
import torch.nn.functional as F

from . import _lib as L
from .models.cheby import ChebyKANLayer
from .models.effkan import KANLinear
from .models.fastkan import FastKANLayer
from .models.nfkan import NaiveFourierKANLayer
from .models.sinekan import SineKANLayer
from .ops import KanSpec


def spec_args(m):
    """(family, in, out, kwargs for KanSpec) of one layer module."""
    if isinstance(m, nn.Linear):
        return L.KV_LINEAR, m.in_features, m.out_features, {}
    if isinstance(m, KANLinear):
        if not isinstance(m.base_activation, nn.SiLU):
            raise L.KanVitError("the fused kernel implements base_activation=SiLU only")
        return L.KV_EFFKAN, m.in_features, m.out_features, dict(grid_size=m.grid_size,
                                                                  spline_order=m.spline_order)
    if isinstance(m, ChebyKANLayer):
        return L.KV_CHEBY, m.inputdim, m.outdim, dict(grid_size=m.degree)
    if isinstance(m, FastKANLayer):
        if m.use_base_update and m.base_activation is not F.silu:
            raise L.KanVitError("grouped FastKAN layers need base_activation=F.silu")
        flags = 0 if m.use_base_update else L.KV_FLAG_NO_BASE
        return L.KV_FASTKAN, m.input_dim, m.output_dim, dict(grid_size=m.num_grids, flags=flags,
                                                             rbf_denom=m.rbf.denominator)
    if isinstance(m, SineKANLayer):
        return L.KV_SINE, m.input_dim, m.output_dim, dict(grid_size=m.grid_size)
    if isinstance(m, NaiveFourierKANLayer):
        return L.KV_FOURIER, m.inputdim, m.outdim, dict(grid_size=m.gridsize)
    raise L.KanVitError(f"no kernel for module type {type(m).__name__}")


def slots_of(m):
    if isinstance(m, nn.Linear):
        return [m.weight, m.bias]
    if isinstance(m, FastKANLayer):
        return m._slots(m.use_base_update)
    return m._slots()


def make_spec(modules, x_share=1):
    """One KanSpec for a list of equal-shaped modules (n_groups = len(modules))."""
    fam, i, o, kw = spec_args(modules[0])
    for m in modules[1:]:
        if spec_args(m) != (fam, i, o, kw):
            raise L.KanVitError("grouped layers must share family, sizes and grid")
    spec = KanSpec(fam, i, o, n_groups=len(modules), x_share=x_share, **kw)
    if fam == L.KV_EFFKAN:
        mods = list(modules)


        def uniform_flag():
            # one uniform knot vector shared by every feature of every grouped module
            if not all(m.grid_is_uniform() for m in mods):
                return 0
            ends = {m.grid_ends() for m in mods}
            return L.KV_FLAG_UNIFORM_GRID if len(ends) == 1 else 0

        spec.dyn_flags = uniform_flag
    return spec
