"""CPU oracle for the KAN-ViT hot path.  TEST INFRASTRUCTURE ONLY.

This file is a torch-CPU *restatement* of the reference's algorithm for the
forward pass of the ViKANformer hot path (the backward is torch autograd
through these functions, which is exactly how the reference obtains it: none
of the reference modules defines a custom backward).

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs may import this package.  The
product (``kan_vit_b200``) never does: it raises when the CUDA library is
missing.

Parity status: PINNED.  ``oracle/make_golden.py`` imports the unmodified
reference from ``/root/reference`` in the build container, runs it on seeded
inputs and commits inputs / weights / outputs / gradients under
``tests/golden/``; ``tests/test_oracle_golden.py`` checks every function here
against those vectors.  (The reference's own tests hold no golden vectors for
this path -- SURVEY.md §8c.)

All functions are *functional*: they take tensors (named like the reference's
``state_dict`` entries) so that autograd gives gradients w.r.t. any of them.
Citations are ``file:line`` under ``/root/reference``.
"""
from __future__ import annotations

import math
from typing import Dict, Optional

import numpy as np
import torch
import torch.nn.functional as F

Tensor = torch.Tensor

# --------------------------------------------------------------------------
# efficient-kan  (models/effkan.py)
# --------------------------------------------------------------------------


def effkan_default_grid(in_features: int, grid_size: int = 5, spline_order: int = 3,
                        grid_range=(-1.0, 1.0), dtype=torch.float32) -> Tensor:
    """Knot vector buffer ``grid (in, G+2k+1)``; models/effkan.py:45-54."""
    h = (grid_range[1] - grid_range[0]) / grid_size
    knots = torch.arange(-spline_order, grid_size + spline_order + 1) * h + grid_range[0]
    return knots.to(dtype).repeat(in_features, 1)


def b_splines(x: Tensor, grid: Tensor, spline_order: int) -> Tensor:
    """Cox-de Boor B-spline bases; models/effkan.py:99-132.

    x (T,in), grid (in, G+2k+1) -> (T, in, G+k).  Order-0 indicator is the
    half-open test ``grid[j] <= x < grid[j+1]`` (effkan.py:115); each level
    p=1..k blends neighbours with the per-feature knot differences
    (effkan.py:116-125).
    """
    assert x.dim() == 2 and x.size(1) == grid.size(0)
    xe = x[:, :, None]
    g = grid[None, :, :]
    basis = torch.logical_and(xe >= g[..., :-1], xe < g[..., 1:]).to(x.dtype)
    for p in range(1, spline_order + 1):
        n = basis.size(-1) - 1  # number of order-p functions
        left_num = xe - g[..., :n]
        left_den = g[..., p:p + n] - g[..., :n]
        right_num = g[..., p + 1:p + 1 + n] - xe
        right_den = g[..., p + 1:p + 1 + n] - g[..., 1:1 + n]
        basis = left_num / left_den * basis[..., :n] + right_num / right_den * basis[..., 1:1 + n]
    return basis


def effkan_forward(x: Tensor, base_weight: Tensor, spline_weight: Tensor,
                   spline_scaler: Optional[Tensor], grid: Tensor, spline_order: int = 3) -> Tensor:
    """KANLinear.forward; models/effkan.py:166-187.

    Y = silu(X) W_b^T + B(X) (W_s * s)^T, K index of the spline GEMM = i*(G+k)+g.
    """
    out_features, in_features = base_weight.shape
    assert x.size(-1) == in_features
    lead = x.shape[:-1]
    x2 = x.reshape(-1, in_features)
    w = spline_weight if spline_scaler is None else spline_weight * spline_scaler[..., None]
    y = F.linear(F.silu(x2), base_weight)
    y = y + F.linear(b_splines(x2, grid, spline_order).reshape(x2.size(0), -1),
                     w.reshape(out_features, -1))
    return y.reshape(*lead, out_features)


# --------------------------------------------------------------------------
# ChebyKAN  (models/cheby.py)
# --------------------------------------------------------------------------


def cheby_forward(x: Tensor, cheby_coeffs: Tensor) -> Tensor:
    """ChebyKANLayer.forward; models/cheby.py:36-48.

    cheby_coeffs is (in, out, deg+1).  T_n(tanh x) is evaluated the way the
    reference does, cos(n*acos(tanh x)), and the output is ALWAYS 2-D
    (cheby.py:38,47).
    """
    in_dim, out_dim, m = cheby_coeffs.shape
    t = torch.tanh(x).reshape(-1, in_dim, 1)
    n = torch.arange(m, device=x.device)
    basis = torch.cos(torch.acos(t) * n)  # (T,in,deg+1)
    return torch.einsum("tin,ion->to", basis, cheby_coeffs).reshape(-1, out_dim)


# --------------------------------------------------------------------------
# FastKAN  (models/fastkan.py)
# --------------------------------------------------------------------------


def fastkan_forward(x: Tensor, ln_weight: Tensor, ln_bias: Tensor, rbf_grid: Tensor,
                    spline_weight: Tensor, base_weight: Optional[Tensor], base_bias: Optional[Tensor],
                    denominator: Optional[float] = None, time_benchmark: bool = False) -> Tensor:
    """FastKANLayer.forward; models/fastkan.py:15-30,66-76.

    RBF branch sees LayerNorm(x) (eps 1e-5), SiLU branch sees raw x
    (fastkan.py:68,74).  spline_weight is (out, in*num_grids), K index i*G+g.
    """
    num_grids = rbf_grid.numel()
    if denominator is None:
        denominator = float(rbf_grid[-1].detach() - rbf_grid[0].detach()) / (num_grids - 1)
    xin = x if time_benchmark else F.layer_norm(x, (x.size(-1),), ln_weight, ln_bias, 1e-5)
    rbf = torch.exp(-(((xin[..., None] - rbf_grid) / denominator) ** 2))
    y = F.linear(rbf.flatten(-2), spline_weight)
    if base_weight is not None:
        y = y + F.linear(F.silu(x), base_weight, base_bias)
    return y


# --------------------------------------------------------------------------
# SineKAN  (models/sinekan.py)
# --------------------------------------------------------------------------

_SINE_A, _SINE_K, _SINE_C = 0.9724108095811765, 0.9884401790754128, 0.999449553483052


def sinekan_phase(input_dim: int, grid_size: int) -> Tensor:
    """Fixed ``phase (1,1,in,G)`` buffer; models/sinekan.py:47,59-75 (+ forward_step :7-23)."""
    grid_phase = torch.arange(1, grid_size + 1).reshape(1, 1, 1, grid_size) / (grid_size + 1)
    input_phase = torch.linspace(0, math.pi, input_dim).reshape(1, 1, input_dim, 1)
    phase = grid_phase + input_phase
    for j in range(1, grid_size):
        phase = (_SINE_A * j ** (-_SINE_K) + _SINE_C) * phase
    return phase


def sinekan_default_freq(grid_size: int, is_first: bool = False, norm_freq: bool = True) -> Tensor:
    """``freq (1,1,1,G)`` init; models/sinekan.py:65-70."""
    f = torch.arange(1, grid_size + 1).float().reshape(1, 1, 1, grid_size)
    return f / (grid_size + 1) ** (1 - is_first) if norm_freq else f


def sinekan_forward(x: Tensor, amplitudes: Tensor, freq: Tensor, phase: Tensor,
                    bias: Optional[Tensor]) -> Tensor:
    """SineKANLayer.forward; models/sinekan.py:81-91.  amplitudes (out,in,G)."""
    out_dim, in_dim, _ = amplitudes.shape
    lead = x.shape[:-1]
    x2 = x.reshape(-1, 1, in_dim, 1)
    s = torch.sin(x2 * freq + phase)[:, 0]  # (T,in,G)
    y = torch.einsum("tig,oig->to", s, amplitudes)
    if bias is not None:
        y = y + bias
    return y.reshape(*lead, out_dim)


# --------------------------------------------------------------------------
# Naive Fourier KAN  (models/nfkan.py)
# --------------------------------------------------------------------------


def fourier_forward(x: Tensor, fouriercoeffs: Tensor, bias: Optional[Tensor]) -> Tensor:
    """NaiveFourierKANLayer.forward; models/nfkan.py:36-52.

    fouriercoeffs (2,out,in,G): [0] multiplies cos(kx), [1] sin(kx), k=1..G.
    Written as the equivalent contraction (the reference materialises
    (T,out,in,G); FourierKAN/fftKAN.py:50-62 documents this einsum form).
    """
    _, out_dim, in_dim, g = fouriercoeffs.shape
    lead = x.shape[:-1]
    x2 = x.reshape(-1, in_dim, 1)
    k = torch.arange(1, g + 1, device=x.device).reshape(1, 1, g)
    kx = k * x2
    y = torch.einsum("tig,oig->to", torch.cos(kx), fouriercoeffs[0])
    y = y + torch.einsum("tig,oig->to", torch.sin(kx), fouriercoeffs[1])
    if bias is not None:
        y = y + bias
    return y.reshape(*lead, out_dim)


# --------------------------------------------------------------------------
# Layer dispatch on a state-dict prefix (shared by MSA / ViT below)
# --------------------------------------------------------------------------


def layer_forward(kind: str, x: Tensor, sd: Dict[str, Tensor], prefix: str) -> Tensor:
    """Apply the reference module of type ``kind`` whose parameters live at ``prefix``.

    ``kind`` uses the reference's ``type=`` strings (model.py:67-80, attention.py:136-170).
    """
    p = prefix
    if kind in ("vanilla", "linear"):
        return F.linear(x, sd[p + "weight"], sd.get(p + "bias"))
    if kind == "efficientkan":
        return effkan_forward(x, sd[p + "base_weight"], sd[p + "spline_weight"],
                              sd.get(p + "spline_scaler"), sd[p + "grid"])
    if kind == "cheby":
        return cheby_forward(x, sd[p + "cheby_coeffs"])
    if kind == "fast":
        return fastkan_forward(x, sd[p + "layernorm.weight"], sd[p + "layernorm.bias"],
                               sd[p + "rbf.grid"], sd[p + "spline_linear.weight"],
                               sd.get(p + "base_linear.weight"), sd.get(p + "base_linear.bias"))
    if kind == "sine":
        return sinekan_forward(x, sd[p + "amplitudes"], sd[p + "freq"], sd[p + "phase"],
                               sd.get(p + "bias"))
    if kind == "fourier":
        return fourier_forward(x, sd[p + "fouriercoeffs"], sd.get(p + "bias"))
    raise ValueError(f"Unknown layer kind: {kind}")


# --------------------------------------------------------------------------
# Attention  (attention.py:112-202, utils.py:137-295)
# --------------------------------------------------------------------------


def attention_core(q: Tensor, k: Tensor, v: Tensor) -> Tensor:
    """softmax(q k^T / sqrt(d_h)) v on (..., N, d_h); attention.py:199-200."""
    s = q @ k.transpose(-1, -2) / (q.size(-1) ** 0.5)
    return torch.softmax(s, dim=-1) @ v


def msa_forward(x: Tensor, sd: Dict[str, Tensor], prefix: str, n_heads: int, kind: str) -> Tensor:
    """MSA.forward; attention.py:181-202.

    Head h sees channels [h*d_h,(h+1)*d_h) and maps d_h->d_h with its own
    q/k/v module; heads are concatenated in order (hstack), no out-projection.
    The reference loops over samples; every per-head map is row-wise, so the
    batch is applied at once here (same function).  For ``fourier`` the
    reference uses Linear maps (attention.py:137).
    """
    b, n, d = x.shape
    dh = d // n_heads
    map_kind = "vanilla" if kind in ("vanilla", "flash-attn", "fourier") else kind
    outs = []
    for h in range(n_heads):
        xs = x[:, :, h * dh:(h + 1) * dh]
        q = layer_forward(map_kind, xs, sd, f"{prefix}q_mappings.{h}.").reshape(b, n, dh)
        k = layer_forward(map_kind, xs, sd, f"{prefix}k_mappings.{h}.").reshape(b, n, dh)
        v = layer_forward(map_kind, xs, sd, f"{prefix}v_mappings.{h}.").reshape(b, n, dh)
        outs.append(attention_core(q, k, v))
    return torch.cat(outs, dim=-1)


def block_forward(x: Tensor, sd: Dict[str, Tensor], prefix: str, n_heads: int, kind: str,
                  kan_mlp: bool = False) -> Tensor:
    """TransformerBlock.forward; model.py:31-37.

    ``kan_mlp=True`` is the cfg-4 extension: ``ff`` is two
    reference KANLinear layers (``ff.0``, ``ff.1``) instead of Linear/ReLU/Linear.
    """
    d = x.size(-1)
    h = F.layer_norm(x, (d,), sd[prefix + "norm1.weight"], sd[prefix + "norm1.bias"], 1e-5)
    x = x + msa_forward(h, sd, prefix + "attn.", n_heads, kind)
    h = F.layer_norm(x, (d,), sd[prefix + "norm2.weight"], sd[prefix + "norm2.bias"], 1e-5)
    if kan_mlp:
        h = layer_forward("efficientkan", h, sd, prefix + "ff.0.")
        h = layer_forward("efficientkan", h, sd, prefix + "ff.1.")
    else:
        h = F.linear(h, sd[prefix + "ff.0.weight"], sd[prefix + "ff.0.bias"])
        h = F.relu(h)
        h = F.linear(h, sd[prefix + "ff.2.weight"], sd[prefix + "ff.2.bias"])
    return x + h


# --------------------------------------------------------------------------
# ViT assembly  (model.py:40-169)
# --------------------------------------------------------------------------


def patchify(images: Tensor, n_patches: int) -> Tensor:
    """VisionTransformer.patchify; model.py:111-126.

    patches[b, i*np+j, c*ph*pw + y*pw + x] = images[b, c, i*ph+y, j*pw+x]; the
    result is float32 regardless of the input dtype (model.py:116).
    """
    b, c, h, w = images.shape
    ph, pw = h // n_patches, w // n_patches
    t = images.reshape(b, c, n_patches, ph, n_patches, pw).permute(0, 2, 4, 1, 3, 5)
    return t.reshape(b, n_patches * n_patches, c * ph * pw).to(torch.float32)


def positional_embeddings(seq_length: int, d: int) -> Tensor:
    """VisionTransformer.positional_embeddings; model.py:128-140.

    pos[i,j] = sin(i / 10000**(j/d)) for even j, cos(same argument) for odd j
    (same exponent j/d for both parities), computed in float64, stored float32.
    """
    i = np.arange(seq_length, dtype=np.float64)[:, None]
    j = np.arange(d, dtype=np.float64)[None, :]
    arg = i / np.power(10000.0, j / d)
    table = np.where((np.arange(d) % 2 == 0)[None, :], np.sin(arg), np.cos(arg))
    return torch.from_numpy(table).to(torch.float32)


def vit_forward(images: Tensor, sd: Dict[str, Tensor], *, n_patches: int, n_blocks: int,
                n_heads: int, kind: str, kan_mlp: bool = False) -> Tensor:
    """VisionTransformer.forward; model.py:142-169.

    ``kind == 'cheby'`` includes the labelled shim of SURVEY §0.3: the
    reference's ChebyKANLayer flattens to 2-D so model.py:150 crashes; the
    patch-embed output is reshaped back to (B, P^2, d).
    """
    patches = patchify(images, n_patches).to(images.dtype)
    b, p2, _ = patches.shape
    tok = layer_forward(kind if kind != "flash-attn" else "vanilla", patches, sd, "linear_mapper.")
    tok = tok.reshape(b, p2, -1)  # cheby shim (no-op for the other kinds)
    d = tok.size(-1)
    cls = sd["v_class"][None].expand(b, -1, -1)
    tok = torch.cat((cls, tok), dim=1)
    pos = sd["pos_embeddings"] if "pos_embeddings" in sd else positional_embeddings(p2 + 1, d)
    x = tok + pos[: tok.size(1)].to(tok.dtype)
    for blk in range(n_blocks):
        if kind == "flash-attn":   # bare attention modules, model.py:93-95,156-159
            x = flash_attention_forward(x, sd, f"blocks.{blk}.", n_heads)
        else:
            x = block_forward(x, sd, f"blocks.{blk}.", n_heads, kind, kan_mlp)
    c = F.layer_norm(x[:, 0], (d,), sd["mlp_head.0.weight"], sd["mlp_head.0.bias"], 1e-5)
    return F.linear(c, sd["mlp_head.1.weight"], sd["mlp_head.1.bias"])


# --------------------------------------------------------------------------
# The two ops either side of the path in the training loop (train.py:22-23,35-40):
# nn.CrossEntropyLoss() and torch.optim.Adam(lr).  The reference calls torch for both, so
# these numpy restatements are pinned against torch itself (tests/test_oracle_golden.py).
# --------------------------------------------------------------------------


def cross_entropy_np(logits: np.ndarray, labels: np.ndarray, ignore_index: int = -100):
    """(loss, d loss / d logits) of nn.CrossEntropyLoss() -- train.py:22,36: class-index targets, mean over
    the rows whose label is not ``ignore_index``."""
    z = logits.astype(np.float64)
    valid = labels != ignore_index
    n = int(valid.sum())
    mx = z.max(axis=1, keepdims=True)
    e = np.exp(z - mx)
    s = e.sum(axis=1, keepdims=True)
    lse = np.log(s[:, 0]) + mx[:, 0]
    rows = np.arange(z.shape[0])
    picked = z[rows, np.where(valid, labels, 0)]
    loss = float(((lse - picked) * valid).sum() / n) if n else float("nan")
    p = e / s
    p[rows, np.where(valid, labels, 0)] -= 1.0
    dz = p * valid[:, None] / max(n, 1)
    return np.float32(loss), dz.astype(np.float32)


def adam_step_np(p: np.ndarray, g: np.ndarray, m: np.ndarray, v: np.ndarray, t: int, lr: float = 1e-3,
                 beta1: float = 0.9, beta2: float = 0.999, eps: float = 1e-8):
    """One torch.optim.Adam update (train.py:23,40; no weight decay / amsgrad), step count ``t`` >= 1.
    fp32 state like torch; returns the new (p, m, v)."""
    f = np.float32
    m = (f(beta1) * m + f(1.0 - beta1) * g).astype(f)
    v = (f(beta2) * v + f(1.0 - beta2) * g * g).astype(f)
    step_size = f(lr / (1.0 - beta1 ** t))
    bc2_sqrt = f(math.sqrt(1.0 - beta2 ** t))
    denom = (np.sqrt(v) / bc2_sqrt + f(eps)).astype(f)
    return (p - step_size * (m / denom)).astype(f), m, v


# --------------------------------------------------------------------------
# FlashAttention module / type="flash-attn"  (attention.py:13-109, utils.py:134-295, model.py:93-95)
# --------------------------------------------------------------------------

_FLASH_EPS = 1e-10  # utils.py:10


def _flash_causal_hidden(nq: int, nk: int) -> Tensor:
    """[nq, nk] bool, True where the reference's causal fill hides key j from query i: the union over tiles of
    ``ones.triu(q_start - k_start + 1)`` with ``q_start = tile_start - max(nk - nq, 0)`` (utils.py:142,175,
    192-194), i.e. j > i - max(nk - nq, 0)."""
    diff = max(nk - nq, 0)
    i = torch.arange(nq)[:, None]
    j = torch.arange(nk)[None, :]
    return j > i - diff


class FlashAttentionFunctionOracle(torch.autograd.Function):
    """utils.py:137-295 with one tile per direction (the bucket sizes only reorder the sums).  The reference
    defines its own backward, so the restatement does too: for rows whose keys are all hidden the custom
    backward is NOT the derivative of the forward, and the kernels follow the reference there as well."""

    @staticmethod
    def forward(ctx, q, k, v, mask, causal):
        neg = -torch.finfo(q.dtype).max
        if mask is not None and mask.ndim == 2:
            mask = mask[:, None, None, :]                        # 'b n -> b 1 1 n', utils.py:154-155
        s = torch.einsum("bhid,bhjd->bhij", q, k) * (q.shape[-1] ** -0.5)
        if mask is not None:
            s = s.masked_fill(~mask, neg)                        # utils.py:186-187
        if causal:
            s = s.masked_fill(_flash_causal_hidden(q.shape[-2], k.shape[-2]), neg)   # utils.py:189-195
        m = s.amax(dim=-1, keepdim=True)
        e = torch.exp(s - m)
        if mask is not None:
            e = e.masked_fill(~mask, 0.0)                        # utils.py:202-203
        rs = e.sum(dim=-1, keepdim=True).clamp(min=_FLASH_EPS)   # utils.py:205-206
        o = torch.einsum("bhij,bhjd->bhid", e, v) / rs
        lse = rs.log() + m                                       # utils.py:222
        ctx.causal, ctx.mask = causal, mask
        ctx.save_for_backward(q, k, v, o, lse)
        return o

    @staticmethod
    def backward(ctx, do):
        q, k, v, o, lse = ctx.saved_tensors
        scale = q.shape[-1] ** -0.5
        s = torch.einsum("bhid,bhjd->bhij", q, k) * scale
        if ctx.causal:                                           # only the causal fill here, utils.py:267-273
            s = s.masked_fill(_flash_causal_hidden(q.shape[-2], k.shape[-2]), -torch.finfo(q.dtype).max)
        p = torch.exp(s - lse)
        if ctx.mask is not None:
            p = p.masked_fill(~ctx.mask, 0.0)                    # utils.py:276-277
        dv = torch.einsum("bhij,bhid->bhjd", p, do)
        dp = torch.einsum("bhid,bhjd->bhij", do, v)
        D = (do * o).sum(dim=-1, keepdim=True)
        ds = p * scale * (dp - D)
        dq = torch.einsum("bhij,bhjd->bhid", ds, k)
        dk = torch.einsum("bhij,bhid->bhjd", ds, q)
        return dq, dk, dv, None, None


def flash_attention_forward(x: Tensor, sd: Dict[str, Tensor], prefix: str, heads: int, *, context=None,
                            mask=None, causal: bool = False) -> Tensor:
    """FlashAttention.forward (attention.py:61-109): bias-free to_q / to_kv / to_out around the function above."""
    ctx = x if context is None else context
    q = F.linear(x, sd[prefix + "to_q.weight"])
    k, v = F.linear(ctx, sd[prefix + "to_kv.weight"]).chunk(2, dim=-1)

    def split(t):
        b, n, hd = t.shape
        return t.view(b, n, heads, hd // heads).permute(0, 2, 1, 3)

    o = FlashAttentionFunctionOracle.apply(split(q), split(k), split(v), mask, causal)
    b, h, n, d = o.shape
    return F.linear(o.permute(0, 2, 1, 3).reshape(b, n, h * d), sd[prefix + "to_out.weight"])


# --------------------------------------------------------------------------
# Harness-side initial state for the timed CPU arm (bench.py --impl reference, cpu_baseline)
# --------------------------------------------------------------------------


def init_state(chw, n_patches: int, n_blocks: int, d: int, n_heads: int, out_d: int, kind: str,
               kan_mlp: bool = False, seed: int = 0) -> Dict[str, Tensor]:
    """A state dict with the reference's key names and shapes (model.py:49-109, attention.py:136-170, the
    constructors in models/*.py) for one VisionTransformer configuration, drawn here so that the CPU arm of
    bench.py does not import the product package.  The distributions follow the reference's scale
    (kaiming-uniform-like Linear / base weights, small spline coefficients, SineKAN / FastKAN defaults); the
    timing of a training step does not depend on the draw.  ``kan_mlp``: ff = (KANLinear, KANLinear) under
    ff.0 / ff.1."""
    g = torch.Generator().manual_seed(seed)
    sd: Dict[str, Tensor] = {}

    def uni(shape, fan_in):
        b = 1.0 / math.sqrt(fan_in)
        return (torch.rand(shape, generator=g) * 2 - 1) * b

    def layer(prefix, k, i, o, grid=None):
        if k in ("vanilla", "flash-attn", "fourier", "linear"):
            sd[prefix + "weight"] = uni((o, i), i)
            sd[prefix + "bias"] = uni((o,), i)
        elif k == "efficientkan":
            sd[prefix + "base_weight"] = uni((o, i), i)
            sd[prefix + "spline_weight"] = torch.randn((o, i, 8), generator=g) * 0.02
            sd[prefix + "spline_scaler"] = uni((o, i), i)
            sd[prefix + "grid"] = effkan_default_grid(i)
        elif k == "cheby":
            deg = 4
            sd[prefix + "cheby_coeffs"] = torch.randn((i, o, deg + 1), generator=g) / (i * (deg + 1))
            sd[prefix + "arange"] = torch.arange(0, deg + 1)
        elif k == "fast":
            sd[prefix + "layernorm.weight"] = torch.ones(i)
            sd[prefix + "layernorm.bias"] = torch.zeros(i)
            sd[prefix + "rbf.grid"] = torch.linspace(-2.0, 2.0, 8)
            sd[prefix + "spline_linear.weight"] = torch.randn((o, i * 8), generator=g).clamp(-2, 2) * 0.1
            sd[prefix + "base_linear.weight"] = uni((o, i), i)
            sd[prefix + "base_linear.bias"] = uni((o,), i)
        elif k == "sine":
            G = grid or 5
            sd[prefix + "amplitudes"] = torch.randn((o, i, G), generator=g) * 0.4 / o / (G / 5.0) ** 0.5
            sd[prefix + "freq"] = sinekan_default_freq(G)
            sd[prefix + "bias"] = torch.ones(1, o) / o
            sd[prefix + "phase"] = sinekan_phase(i, G)
        else:
            raise ValueError(f"Unknown layer kind: {k}")

    c, hh, ww = chw
    input_d = c * (hh // n_patches) * (ww // n_patches)
    layer("linear_mapper.", kind, input_d, d, grid=28)
    sd["v_class"] = torch.randn((1, d), generator=g)
    dh = d // n_heads
    map_kind = "vanilla" if kind in ("vanilla", "flash-attn", "fourier") else kind
    for b in range(n_blocks):
        p = f"blocks.{b}."
        for n in ("norm1", "norm2"):
            sd[p + n + ".weight"] = torch.ones(d)
            sd[p + n + ".bias"] = torch.zeros(d)
        for nm in ("q", "k", "v"):
            for h in range(n_heads):
                layer(f"{p}attn.{nm}_mappings.{h}.", map_kind, dh, dh, grid=4)
        if kan_mlp:
            layer(p + "ff.0.", "efficientkan", d, 4 * d)
            layer(p + "ff.1.", "efficientkan", 4 * d, d)
        else:
            layer(p + "ff.0.", "linear", d, 4 * d)
            layer(p + "ff.2.", "linear", 4 * d, d)
    sd["mlp_head.0.weight"] = torch.ones(d)
    sd["mlp_head.0.bias"] = torch.zeros(d)
    layer("mlp_head.1.", "linear", d, out_d)
    return sd
