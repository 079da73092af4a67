"""Seeded, bf16-representable weights and inputs for the parity cases at BASELINE geometry.  TEST INFRASTRUCTURE.

The large cases (768 <-> 3072 KANLinear, a d=768 / 12-head / 197-token block, cfg3 ViTs) cannot commit their
weights as fixtures (a KANLinear(768, 3072) alone is 85 MB), so weights and inputs are REGENERATED from a seed
with the CPU generator, which is bit-reproducible across machines for one torch build.  The same functions are
used by oracle/make_golden.py --big (which runs the unmodified reference on them and commits sub-sampled
outputs / gradients) and by tests/test_gpu_big_parity.py.

Every value is rounded to bf16 (the bf16 path is compared with the fp32 reference evaluated
on the bf16-rounded inputs and weights); `spline_scaler` values are powers of two so that the effective weight
spline_weight * spline_scaler (models/effkan.py:166-172) is bf16-representable too.
"""
import math
import zlib

import torch


def bf16r(t):
    return t.to(torch.bfloat16).to(torch.float32)


def randn(shape, seed, scale=1.0):
    return torch.randn(shape, generator=torch.Generator().manual_seed(seed)) * scale


def seeded_input(shape, seed, scale=1.0):
    return bf16r(randn(shape, seed, scale))


_KEEP = ("grid", "phase", "arange", "pos_embeddings", "freq")


def seeded_state(state_dict, seed):
    """A new state dict with the shapes / keys of ``state_dict`` (reference module or this repo's module -- the key
    lists are identical) and seeded values.  Buffers that define the basis (knots, phases, RBF centres, SineKAN
    frequencies) keep the constructor's values."""
    out = {}
    for k in sorted(state_dict):
        v = state_dict[k]
        leaf = k.rsplit(".", 1)[-1]
        if not v.is_floating_point() or leaf in _KEEP:
            out[k] = v.clone()
            continue
        s = (seed * 1000003 + zlib.crc32(k.encode())) % (2 ** 31)
        g = torch.Generator().manual_seed(s)
        if leaf == "spline_scaler":
            out[k] = 2.0 ** torch.randint(-1, 2, v.shape, generator=g).to(torch.float32)
            continue
        if leaf == "weight" and v.dim() == 1:           # LayerNorm gamma (the only 1-D weights on the path)
            out[k] = bf16r(1.0 + 0.1 * torch.randn(v.shape, generator=g))
            continue
        if v.dim() == 1 or leaf == "bias":
            out[k] = bf16r(0.1 * torch.randn(v.shape, generator=g))
            continue
        if leaf == "v_class":
            out[k] = bf16r(torch.randn(v.shape, generator=g))
            continue
        if leaf == "cheby_coeffs":                      # (in, out, deg+1)
            fan = v.shape[0] * v.shape[2]
        elif leaf == "fouriercoeffs":                   # (2, out, in, G)
            fan = 2 * v.shape[2] * v.shape[3]
        elif leaf == "amplitudes":                      # (out, in, G)
            fan = v.shape[1] * v.shape[2]
        else:                                           # (out, in[, M]) and (out, in * M)
            fan = v.shape[1]
        out[k] = bf16r(torch.randn(v.shape, generator=g) / math.sqrt(fan))
    return out


def sample_stride(numel, target=4096):
    """Odd stride (coprime with the power-of-two feature widths) that keeps about ``target`` elements."""
    k = max(1, numel // target)
    return k | 1


def subsample(t, target=4096):
    k = sample_stride(t.numel(), target)
    return t.detach().reshape(-1)[::k].clone(), k
