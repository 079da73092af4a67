"""torch.autograd.Function wrappers around the C ABI (include/kanvit_b200.h).

PyTorch is plumbing here: it owns the device memory, the stream and the autograd graph.
All arithmetic of the path runs in libkanvit_b200.so.  No CPU fallback.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence

import torch

from . import _lib as L

_FAMILY_GRAD_SLOTS = {
    L.KV_LINEAR: (0, 1),
    L.KV_EFFKAN: (0, 1, 2),
    L.KV_CHEBY: (0,),
    L.KV_FASTKAN: (0, 1, 3, 4, 5),
    L.KV_SINE: (0, 1, 3),
    L.KV_FOURIER: (0, 1),
}
_FAMILY_BIAS_SLOT = {L.KV_LINEAR: 1, L.KV_FASTKAN: 5, L.KV_SINE: 3, L.KV_FOURIER: 1}


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _require_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise L.KanVitError("kan_vit_b200 ops need CUDA tensors (there is no CPU path)")


# Optional kernel timing hook (bench.py): when PROFILE is a dict, calls whose name it selects
# are bracketed by CUDA events on the launching stream; entries are (start, end, work) with
# work = algorithmic FLOPs (GEMM-like kernels) or bytes (memory-bound kernels) of that call.
PROFILE = None
PROFILE_ONLY = None  # None = every call, else a set of names

# Optional gradient sink (set by trainer.TrainStep around backward): an object with
#   target(param) -> the tensor the gradient of `param` must be ADDED to (its .grad view in the flat bucket
#                    buffer), or None when the parameter is not managed;
#   done(param)   -> called once that gradient has been written (bucket bookkeeping / all-reduce trigger).
# With a sink, kv_kan_unpack_grads accumulates straight into those buffers and autograd gets no parameter
# gradients back, which removes one small `add` launch per parameter tensor per step.
GRAD_SINK = None

# Arithmetic of the KAN / Linear GEMMs: "fp32" = exact SIMT kernels (parity 1e-5), "bf16" = tcgen05
# tensor-core kernels with bf16 operands and fp32 accumulation in TMEM (parity 2e-2).  Activations
# stay fp32 in HBM in both modes.
_PRECISION = "fp32"
_DEBUG_BITS = int(__import__("os").environ.get("KV_DEBUG_BITS", "0"))  # kernel timing experiments only
_NO_FUSED_ATTN = __import__("os").environ.get("KV_NO_FUSED_ATTN") == "1"  # A/B experiments: mma.sync attention route


def set_precision(mode: str):
    """Select the GEMM arithmetic for subsequent forward passes: "fp32" or "bf16"."""
    global _PRECISION
    if mode not in ("fp32", "bf16"):
        raise ValueError(f"precision must be 'fp32' or 'bf16', not {mode!r}")
    _PRECISION = mode


def get_precision() -> str:
    return _PRECISION


class precision:
    """Context manager: ``with kan_vit_b200.precision("bf16"): ...``"""

    def __init__(self, mode):
        self.mode = mode

    def __enter__(self):
        self.prev = _PRECISION
        set_precision(self.mode)

    def __exit__(self, *exc):
        set_precision(self.prev)


def _call(name, fn, *args, work=0.0, tag=None):
    if PROFILE is None or (PROFILE_ONLY is not None and name not in PROFILE_ONLY):
        L.check(fn(*args), name)
        return
    s = torch.cuda.Event(enable_timing=True)
    e = torch.cuda.Event(enable_timing=True)
    s.record()
    L.check(fn(*args), name)
    e.record()
    PROFILE.setdefault(name, []).append((s, e, work, tag))


def _kan_tag(spec, T):
    return f"T{T} g{spec.n_groups} {spec.in_features}->{spec.out_features} M{spec.M}"


def _kan_flops(spec, T):
    return 2.0 * T * spec.in_features * spec.out_features * spec.M * spec.n_groups


class KanSpec:
    """Static description of one KAN / Linear layer, or of n_groups equal-shaped layers
    applied to column slices of one activation matrix (kv_kan_desc_t without T / strides)."""

    def __init__(self, family: int, in_features: int, out_features: int, n_groups: int = 1,
                 x_share: int = 1, grid_size: int = 0, spline_order: int = 0, flags: int = 0,
                 rbf_denom: float = 1.0):
        self.family = family
        self.in_features = in_features
        self.out_features = out_features
        self.n_groups = n_groups
        self.x_share = x_share
        self.grid_size = grid_size
        self.spline_order = spline_order
        self.flags = flags
        self.rbf_denom = float(rbf_denom)
        if family == L.KV_LINEAR:
            self.M = 1
        elif family == L.KV_EFFKAN:
            self.M = grid_size + spline_order + 1
        elif family == L.KV_CHEBY:
            self.M = grid_size + 1
        elif family == L.KV_FASTKAN:
            self.M = grid_size + (0 if flags & L.KV_FLAG_NO_BASE else 1)
        elif family == L.KV_SINE:
            self.M = grid_size
        elif family == L.KV_FOURIER:
            self.M = 2 * grid_size
        else:
            raise ValueError(f"unknown family {family}")
        if self.M > L.KV_MAX_M:
            raise L.KanVitError(f"{self.M} basis functions per input exceed KV_MAX_M={L.KV_MAX_M}")
        d = self.desc(0, 0, 0)
        self.fparam_stride = max(L.load().kv_kan_fparam_count(C.byref(d)), 0)
        self._tbl_cache = None     # (ptrs, device tensor)
        self.dyn_flags = None      # callable -> per-call flags (effkan: KV_FLAG_UNIFORM_GRID when the grid is)
        self._goff_cache = None    # (shapes key, offsets tensor, mask, total, views meta)
        self._sink_tbl = None      # (gradient target pointers, device table) when ops.GRAD_SINK is in use

    @property
    def n_slices(self):
        return self.n_groups // self.x_share

    def tc_sizes(self, d):
        """(forward image, input-gradient image, dY image) bytes of the tensor-core path for this call,
        or None when the layer has no tensor-core kernel (it then runs on the fp32 kernels)."""
        a, b, c = C.c_size_t(0), C.c_size_t(0), C.c_size_t(0)
        rc = L.load().kv_kan_tc_bytes(C.byref(d), C.byref(a), C.byref(b), C.byref(c))
        return (a.value, b.value, c.value) if rc == 0 else None

    def desc(self, T, ldx, ldy, rows_in=0, rows_out=0, row_shift=0, extra_flags=0) -> L.KanDesc:
        return L.KanDesc(self.family, T, self.in_features, self.out_features, self.n_groups,
                         self.x_share, self.M, self.grid_size, self.spline_order, ldx, ldy,
                         rows_in, rows_out, row_shift, self.flags | extra_flags, self.rbf_denom,
                         getattr(self, "fparam_stride", 1 << 30), _DEBUG_BITS)

    # -- parameter pointer table (device, cached while the parameter storages stay put)
    def ptr_table(self, flat: Sequence[Optional[torch.Tensor]], device) -> torch.Tensor:
        ptrs = tuple(0 if t is None else t.data_ptr() for t in flat)
        if self._tbl_cache is None or self._tbl_cache[0] != ptrs or self._tbl_cache[1].device != device:
            tbl = torch.tensor(ptrs, dtype=torch.int64).to(device)
            self._tbl_cache = (ptrs, tbl)
        return self._tbl_cache[1]

    # -- packed weights / prologue parameters (cached until a parameter is modified)
    def packed(self, flat, device, tc_sizes, flags=0):
        # No cache across calls: fused optimizers (torch.optim.Adam(fused=True)) update parameters
        # without bumping Tensor._version, so a (data_ptr, version) key would serve stale weights.
        # The pack is one small memory-bound launch per layer; forward hands it to backward via ctx.
        lib = L.load()
        g, o, K = self.n_groups, self.out_features, self.in_features * self.M
        tbl = self.ptr_table(flat, device)
        # the fp32 pack feeds the exact kernels; a layer whose forward and both gradients run on the
        # tensor cores only needs the bf16 operand images (cut straight from the module tensors)
        all_tc = tc_sizes is not None and tc_sizes[1] > 0 and tc_sizes[2] > 0
        wp = None if all_tc else torch.empty((g, o, K), dtype=torch.float32, device=device)
        fp = torch.empty((g, max(self.fparam_stride, 1)), dtype=torch.float32, device=device)
        has_bias = self.family in _FAMILY_BIAS_SLOT and any(
            flat[q * L.KV_PTRS_PER_GROUP + _FAMILY_BIAS_SLOT[self.family]] is not None for q in range(g))
        bias = torch.empty((g, o), dtype=torch.float32, device=device) if has_bias else None
        d = self.desc(0, self.n_slices * self.in_features, g * o, extra_flags=flags)
        wt = wtt = None
        if tc_sizes is not None:
            wt = torch.empty((tc_sizes[0],), dtype=torch.uint8, device=device)
            if tc_sizes[1]:
                wtt = torch.empty((tc_sizes[1],), dtype=torch.uint8, device=device)
        _call("kv_kan_pack", lib.kv_kan_pack, C.byref(d), tbl.data_ptr(), L.ptr(wp), L.ptr(wt), L.ptr(wtt),
                                fp.data_ptr() if self.fparam_stride else None, L.ptr(bias), _stream())
        pk = {"wp": wp, "fp": fp if self.fparam_stride else None, "bias": bias, "tbl": tbl, "wt": wt, "wtt": wtt}
        return pk

    # -- flat gradient buffer layout (byte offsets per pointer-table slot)
    def grad_layout(self, flat, device):
        key = tuple(None if t is None else tuple(t.shape) for t in flat)
        if self._goff_cache is not None and self._goff_cache[0] == key and self._goff_cache[1].device == device:
            return self._goff_cache[1:]
        offs, meta, total = [], [], 0
        slots = _FAMILY_GRAD_SLOTS[self.family]
        for idx, t in enumerate(flat):
            s = idx % L.KV_PTRS_PER_GROUP
            if t is None or s not in slots or not t.is_floating_point():
                offs.append(-1)
                meta.append(None)
                continue
            offs.append(total * 4)
            meta.append((total, t.numel(), tuple(t.shape)))
            total += (t.numel() + 3) // 4 * 4  # keep every gradient 16-byte aligned
        off_t = torch.tensor(offs, dtype=torch.int64).to(device)
        self._goff_cache = (key, off_t, meta, total)
        return off_t, meta, total


def _kan_param_grads(spec, flat, d, pk, x, dy, ymask, dyimg, mean, rstd, dfp):
    """Weight-gradient GEMM + product-rule scatter back to the module tensors.  Returns the list of gradients for
    autograd (None where ops.GRAD_SINK took them).  dy may be None when only the bf16 dY image exists."""
    lib = L.load()
    dev = x.device
    T = x.shape[0]
    g, o, K = spec.n_groups, spec.out_features, spec.in_features * spec.M
    st = _stream()
    grads: List[Optional[torch.Tensor]] = [None] * len(flat)
    dwp = torch.zeros((g, o, K), dtype=torch.float32, device=dev)
    dbias = torch.zeros((g, o), dtype=torch.float32, device=dev) if pk["bias"] is not None else None
    _call("kv_kan_bwd_weight", lib.kv_kan_bwd_weight, C.byref(d), x.data_ptr(), L.ptr(dy), L.ptr(ymask),
          L.ptr(dyimg), L.ptr(pk["fp"]), L.ptr(mean), L.ptr(rstd), dwp.data_ptr(),
          L.ptr(dbias), st, work=_kan_flops(spec, T), tag=_kan_tag(spec, T))
    off_t, meta, total = spec.grad_layout(flat, dev)
    sink = GRAD_SINK
    wanted = [idx for idx, m in enumerate(meta) if m is not None and flat[idx].requires_grad]
    targets = {idx: sink.target(flat[idx]) for idx in wanted} if sink is not None else {}
    if sink is not None and all(t is not None and t.is_contiguous() for t in targets.values()):
        ptrs = tuple(targets[idx].data_ptr() if idx in targets else 0 for idx in range(len(flat)))
        if spec._sink_tbl is None or spec._sink_tbl[0] != ptrs:
            spec._sink_tbl = (ptrs, torch.tensor(ptrs, dtype=torch.int64).to(dev))
        d2 = L.KanDesc.from_buffer_copy(d)
        d2.flags |= L.KV_FLAG_ACCUM_GRADS
        _call("kv_kan_unpack_grads", lib.kv_kan_unpack_grads, C.byref(d2), dwp.data_ptr(), L.ptr(dfp),
              L.ptr(dbias), pk["tbl"].data_ptr(), spec._sink_tbl[1].data_ptr(), st)
        for idx in wanted:
            sink.done(flat[idx])
    else:
        gflat = torch.empty((total,), dtype=torch.float32, device=dev)
        gtbl = torch.where(off_t >= 0, off_t + gflat.data_ptr(), torch.zeros_like(off_t))
        _call("kv_kan_unpack_grads", lib.kv_kan_unpack_grads, C.byref(d), dwp.data_ptr(), L.ptr(dfp),
              L.ptr(dbias), pk["tbl"].data_ptr(), gtbl.data_ptr(), st)
        for idx, m in enumerate(meta):
            if m is not None and flat[idx].requires_grad:
                grads[idx] = gflat[m[0]:m[0] + m[1]].view(m[2])
    return grads


class _KanFn(torch.autograd.Function):
    """Y = KAN(x) (+bias) (+pos) [relu] (+residual); optional class-token rows."""

    @staticmethod
    def forward(ctx, x, residual, v_class, pos, spec: KanSpec, cfg: dict, *flat):
        lib = L.load()
        _require_cuda(x, residual, v_class, pos, *flat)
        assert x.dim() == 2 and x.dtype == torch.float32 and x.is_contiguous()
        T = x.shape[0]
        n_in = spec.n_slices * spec.in_features
        n_out = spec.n_groups * spec.out_features
        assert x.shape[1] == n_in
        rows_in, rows_out, row_shift = cfg.get("rows_in", 0), cfg.get("rows_out", 0), cfg.get("row_shift", 0)
        rows = T if rows_in == 0 else (T // rows_in) * rows_out
        extra = L.KV_FLAG_RELU if cfg.get("relu") else 0
        if spec.dyn_flags is not None:
            extra |= spec.dyn_flags()
        tc_sizes = None
        if _PRECISION == "bf16":
            d = spec.desc(T, n_in, n_out, rows_in, rows_out, row_shift, extra | L.KV_FLAG_BF16_TC)
            tc_sizes = spec.tc_sizes(d)
        if tc_sizes is None:
            d = spec.desc(T, n_in, n_out, rows_in, rows_out, row_shift, extra)
        pk = spec.packed(flat, x.device, tc_sizes, d.flags & (L.KV_FLAG_UNIFORM_GRID | L.KV_FLAG_BF16_TC))
        pk["tc_sizes"] = tc_sizes
        mean = rstd = None
        if spec.family == L.KV_FASTKAN and not (spec.flags & L.KV_FLAG_NO_LAYERNORM):
            mean = torch.empty((T, spec.n_slices), dtype=torch.float32, device=x.device)
            rstd = torch.empty_like(mean)
            _call("kv_row_stats", lib.kv_row_stats, x.data_ptr(), T, n_in, spec.n_slices, spec.in_features, 1e-5,
                                     mean.data_ptr(), rstd.data_ptr(), _stream())
        y = torch.empty((rows, n_out), dtype=torch.float32, device=x.device)
        if residual is not None:
            assert residual.shape == y.shape and residual.is_contiguous()
        if pos is not None:
            assert pos.is_contiguous() and pos.shape[-1] == n_out
        _call("kv_kan_fwd", lib.kv_kan_fwd, C.byref(d), x.data_ptr(), L.ptr(pk["wp"]), L.ptr(pk["wt"]), L.ptr(pk["fp"]),
                               L.ptr(pk["bias"]), L.ptr(mean), L.ptr(rstd), L.ptr(residual), L.ptr(pos),
                               y.data_ptr(), _stream(), work=_kan_flops(spec, T), tag=_kan_tag(spec, T))
        if v_class is not None:
            assert rows_in > 0 and row_shift == 1 and spec.n_groups == 1
            _call("kv_cls_rows_fwd", lib.kv_cls_rows_fwd, v_class.data_ptr(), L.ptr(pos), y.data_ptr(), T // rows_in,
                                        rows_out, n_out, _stream())
        ctx.spec, ctx.cfg, ctx.flat, ctx.pk = spec, cfg, flat, pk
        ctx.d = d
        ctx.has_res, ctx.has_cls = residual is not None, v_class is not None
        ctx.save_for_backward(x, y if cfg.get("relu") else None, mean, rstd)
        return y

    @staticmethod
    def backward(ctx, dy):
        lib = L.load()
        spec, flat, pk, d = ctx.spec, ctx.flat, ctx.pk, ctx.d
        x, ymask, mean, rstd = ctx.saved_tensors
        dy = dy.contiguous()
        dev = x.device
        T = x.shape[0]
        g, o, K = spec.n_groups, spec.out_features, spec.in_features * spec.M
        st = _stream()
        need_w = any(t is not None and t.requires_grad for t in flat)
        grads: List[Optional[torch.Tensor]] = [None] * len(flat)
        dfp = None
        if spec.family in (L.KV_SINE, L.KV_FASTKAN):
            dfp = torch.zeros((g, spec.fparam_stride), dtype=torch.float32, device=dev)
        dyimg = None
        if pk["tc_sizes"] is not None and pk["tc_sizes"][2]:
            dyimg = torch.empty((pk["tc_sizes"][2],), dtype=torch.uint8, device=dev)
            _call("kv_kan_dy_image", lib.kv_kan_dy_image, C.byref(d), dy.data_ptr(), L.ptr(ymask), dyimg.data_ptr(), st,
                  work=6.0 * T * g * o)
        dx = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty_like(x)
            dxn = None
            if mean is not None:
                dxn = torch.empty((T, g * spec.in_features), dtype=torch.float32, device=dev)
            _call("kv_kan_bwd_input", lib.kv_kan_bwd_input, C.byref(d), x.data_ptr(), dy.data_ptr(), L.ptr(ymask),
                                         L.ptr(pk["wp"]), L.ptr(pk["wtt"]), L.ptr(dyimg), L.ptr(pk["fp"]), L.ptr(mean),
                                         L.ptr(rstd), dx.data_ptr(), L.ptr(dxn), L.ptr(dfp), st, work=_kan_flops(spec, T), tag=_kan_tag(spec, T))
            if dxn is not None:
                _call("kv_fastkan_ln_bwd", lib.kv_fastkan_ln_bwd, C.byref(d), x.data_ptr(), dxn.data_ptr(), pk["fp"].data_ptr(),
                                              mean.data_ptr(), rstd.data_ptr(), dx.data_ptr(),
                                              dfp.data_ptr(), st)
        elif need_w and spec.family in (L.KV_SINE, L.KV_FASTKAN):
            # d freq / d layernorm.{weight,bias} come out of the input-gradient pass
            dx_tmp = torch.empty_like(x)
            dxn = None
            if mean is not None:
                dxn = torch.empty((T, g * spec.in_features), dtype=torch.float32, device=dev)
            _call("kv_kan_bwd_input", lib.kv_kan_bwd_input, C.byref(d), x.data_ptr(), dy.data_ptr(), L.ptr(ymask),
                                         L.ptr(pk["wp"]), L.ptr(pk["wtt"]), L.ptr(dyimg), L.ptr(pk["fp"]), L.ptr(mean),
                                         L.ptr(rstd), dx_tmp.data_ptr(), L.ptr(dxn), L.ptr(dfp), st, work=_kan_flops(spec, T), tag=_kan_tag(spec, T))
            if dxn is not None:
                _call("kv_fastkan_ln_bwd", lib.kv_fastkan_ln_bwd, C.byref(d), x.data_ptr(), dxn.data_ptr(), pk["fp"].data_ptr(),
                                              mean.data_ptr(), rstd.data_ptr(), dx_tmp.data_ptr(),
                                              dfp.data_ptr(), st)
        if need_w:
            grads = _kan_param_grads(spec, flat, d, pk, x, dy, ymask, dyimg, mean, rstd, dfp)
        dres = dy if ctx.has_res else None
        dcls = None
        if ctx.has_cls:
            n_out = g * o
            dcls = torch.empty((1, n_out), dtype=torch.float32, device=dev)
            _call("kv_cls_rows_bwd", lib.kv_cls_rows_bwd, dy.data_ptr(), dcls.data_ptr(), T // d.rows_in, d.rows_out, n_out, st)
        return (dx, dres, dcls, None, None, None) + tuple(grads)


def kan_apply(x2d, spec: KanSpec, group_slots: Sequence[Sequence[Optional[torch.Tensor]]], *,
              relu=False, residual=None, pos=None, v_class=None, rows_in=0, rows_out=0, row_shift=0):
    """Apply the layer(s) described by ``spec``; ``group_slots[g]`` lists the group's tensors in
    pointer-table slot order (see kv_kan_pack in include/kanvit_b200.h)."""
    flat = []
    for slots in group_slots:
        s = list(slots) + [None] * (L.KV_PTRS_PER_GROUP - len(slots))
        flat.extend(s)
    cfg = {"relu": relu, "rows_in": rows_in, "rows_out": rows_out, "row_shift": row_shift}
    if not x2d.is_contiguous():
        x2d = x2d.contiguous()
    if x2d.dtype != torch.float32:
        x2d = x2d.float()
    return _KanFn.apply(x2d, residual, v_class, pos, spec, cfg, *flat)


class _AttnFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, qkv, residual, B, N, H, dh):
        lib = L.load()
        _require_cuda(qkv, residual)
        assert qkv.is_contiguous() and qkv.shape == (B * N, 3 * H * dh)
        out = torch.empty((B * N, H * dh), dtype=torch.float32, device=qkv.device)
        lse = torch.empty((B, H, N), dtype=torch.float32, device=qkv.device)
        # bf16 mode: tensor-core kernel where it serves the head shape, else the exact kernel
        tc = _PRECISION == "bf16" and dh <= 64 and dh % 4 == 0
        if tc:
            # shape probe (B = 0 launches nothing); the backward needs the larger shared-memory tile
            tc = lib.kv_attn_bwd_bf16(None, None, None, None, 0, N, H, dh, None) == 0
        name, fn = ("kv_attn_fwd_bf16", lib.kv_attn_fwd_bf16) if tc else ("kv_attn_fwd", lib.kv_attn_fwd)
        _call(name, fn, qkv.data_ptr(), L.ptr(residual), out.data_ptr(), lse.data_ptr(),
              B, N, H, dh, _stream(), work=4.0 * B * H * N * N * dh)
        ctx.tc = tc
        ctx.save_for_backward(qkv, lse)
        ctx.dims = (B, N, H, dh)
        ctx.has_res = residual is not None
        return out

    @staticmethod
    def backward(ctx, dout):
        lib = L.load()
        qkv, lse = ctx.saved_tensors
        B, N, H, dh = ctx.dims
        dout = dout.contiguous()
        dqkv = torch.empty_like(qkv)
        name, fn = ("kv_attn_bwd_bf16", lib.kv_attn_bwd_bf16) if ctx.tc else ("kv_attn_bwd", lib.kv_attn_bwd)
        _call(name, fn, qkv.data_ptr(), dout.data_ptr(), lse.data_ptr(), dqkv.data_ptr(),
              B, N, H, dh, _stream(), work=8.0 * B * H * N * N * dh)
        return dqkv, (dout if ctx.has_res else None), None, None, None, None


class _QkvAttnFn(torch.autograd.Function):
    """MSA.forward (attention.py:181-202) in bf16 mode as ONE autograd node: the grouped q/k/v KAN layer writes the
    bf16 attention image (tensor-core operand layout) straight from its GEMM epilogue, the tcgen05 attention kernels
    read it with bulk copies, and the attention backward emits the bf16 dY image the layer's gradient kernels read.
    Neither an fp32 qkv / dqkv tensor nor a dy_image pass exists on this route."""

    @staticmethod
    def forward(ctx, x, residual, spec: KanSpec, dims, *flat):
        lib = L.load()
        B, N, H, dh = dims
        _require_cuda(x, residual, *flat)
        assert x.dim() == 2 and x.dtype == torch.float32 and x.is_contiguous()
        T = x.shape[0]
        n_in = spec.n_slices * spec.in_features
        n_out = spec.n_groups * spec.out_features
        extra = L.KV_FLAG_BF16_TC | (spec.dyn_flags() if spec.dyn_flags is not None else 0)
        d = spec.desc(T, n_in, n_out, 0, 0, 0, extra)
        tc_sizes = spec.tc_sizes(d)
        pk = spec.packed(flat, x.device, tc_sizes, d.flags & (L.KV_FLAG_UNIFORM_GRID | L.KV_FLAG_BF16_TC))
        pk["tc_sizes"] = tc_sizes
        image = torch.empty((lib.kv_attn_image_bytes(B, N, H, dh),), dtype=torch.uint8, device=x.device)
        _call("kv_kan_fwd", lib.kv_kan_fwd_attn_image, C.byref(d), x.data_ptr(), pk["wt"].data_ptr(), L.ptr(pk["fp"]),
              L.ptr(pk["bias"]), None, None, image.data_ptr(), N, _stream(), work=_kan_flops(spec, T), tag=_kan_tag(spec, T))
        out = torch.empty((T, H * dh), dtype=torch.float32, device=x.device)
        lse = torch.empty((B, H, N), dtype=torch.float32, device=x.device)
        _call("kv_attn_fwd_bf16", lib.kv_attn_image_fwd, image.data_ptr(), L.ptr(residual), out.data_ptr(), lse.data_ptr(),
              B, N, H, dh, _stream(), work=4.0 * B * H * N * N * dh)
        ctx.spec, ctx.flat, ctx.pk, ctx.d, ctx.dims = spec, flat, pk, d, dims
        ctx.has_res = residual is not None
        ctx.save_for_backward(x, image, lse)
        return out

    @staticmethod
    def backward(ctx, dout):
        lib = L.load()
        spec, flat, pk, d = ctx.spec, ctx.flat, ctx.pk, ctx.d
        B, N, H, dh = ctx.dims
        x, image, lse = ctx.saved_tensors
        dout = dout.contiguous()
        dev, T, st = x.device, x.shape[0], _stream()
        t8, o8 = C.c_int(0), C.c_int(0)
        L.check(lib.kv_kan_dy_image_geom(C.byref(d), C.byref(t8), C.byref(o8)), "kv_kan_dy_image_geom")
        dyimg = torch.empty((pk["tc_sizes"][2],), dtype=torch.uint8, device=dev)
        _call("kv_attn_bwd_bf16", lib.kv_attn_image_bwd, image.data_ptr(), dout.data_ptr(), lse.data_ptr(), dyimg.data_ptr(),
              B, N, H, dh, t8.value, o8.value, st, work=8.0 * B * H * N * N * dh)
        dx = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty_like(x)
            _call("kv_kan_bwd_input", lib.kv_kan_bwd_input, C.byref(d), x.data_ptr(), None, None, None,
                  pk["wtt"].data_ptr(), dyimg.data_ptr(), L.ptr(pk["fp"]), None, None, dx.data_ptr(), None, None, st,
                  work=_kan_flops(spec, T), tag=_kan_tag(spec, T))
        grads = [None] * len(flat)
        if any(t is not None and t.requires_grad for t in flat):
            grads = _kan_param_grads(spec, flat, d, pk, x, None, None, dyimg, None, None, None)
        return (dx, dout if ctx.has_res else None, None, None) + tuple(grads)


def qkv_attention_fused_ok(spec: KanSpec, flat, N, H, dh) -> bool:
    """True when MSA can take the fused image route: bf16 mode, q/k/v maps that merge into one layer per head
    (efficient-kan on its uniform grid, ChebyKAN) without bias, both gradients on the tensor cores, and a head shape
    the tcgen05 attention kernels serve (N <= 256, d_head <= 64, d_head % 8 == 0)."""
    if _PRECISION != "bf16" or spec.x_share != 3 or spec.family not in (L.KV_EFFKAN, L.KV_CHEBY) or _NO_FUSED_ATTN:
        return False
    if spec.out_features != dh or spec.n_groups != 3 * H:
        return False
    lib = L.load()
    if lib.kv_attn_image_fwd(None, None, None, None, 0, N, H, dh, None) != 0:
        return False
    extra = L.KV_FLAG_BF16_TC | (spec.dyn_flags() if spec.dyn_flags is not None else 0)
    if spec.family == L.KV_EFFKAN and not (extra & L.KV_FLAG_UNIFORM_GRID):
        return False
    d = spec.desc(8, spec.n_slices * spec.in_features, spec.n_groups * spec.out_features, 0, 0, 0, extra)
    sz = spec.tc_sizes(d)
    return sz is not None and sz[1] > 0 and sz[2] > 0


def qkv_attention(x2d, spec: KanSpec, group_slots, B, N, H, dh, residual=None):
    """x2d [B*N, H*dh] -> softmax(q k^T / sqrt(dh)) v (+ residual) with q, k, v = the per-head maps of ``spec``."""
    flat = []
    for slots in group_slots:
        flat.extend(list(slots) + [None] * (L.KV_PTRS_PER_GROUP - len(slots)))
    return _QkvAttnFn.apply(x2d.contiguous(), residual, spec, (B, N, H, dh), *flat)


def attention_tc_ok(N, H, dh) -> bool:
    """True when the bf16 tensor-core attention kernels serve this head shape (shape probe, nothing launched)."""
    return dh <= 64 and dh % 4 == 0 and L.load().kv_attn_bwd_bf16(None, None, None, None, 0, N, H, dh, None) == 0


def attention(qkv, B, N, H, dh, residual=None):
    """softmax(q k^T / sqrt(dh)) v per (sample, head) on qkv [B*N, H*3*dh] -> [B*N, H*dh] (+residual)."""
    return _AttnFn.apply(qkv.contiguous(), residual, B, N, H, dh)


class _LayerNormFn(torch.autograd.Function):
    """fork=True also returns x itself (the skip branch of model.py:34,36): autograd then hands the gradients
    of BOTH uses of x to one backward call, and the kernel adds the skip gradient while it writes dX --
    instead of a separate elementwise add over the whole activation."""

    @staticmethod
    def forward(ctx, x, gamma, beta, eps, ldx, fork=False):
        lib = L.load()
        _require_cuda(x, gamma, beta)
        T, d = x.shape
        assert x.stride(1) == 1 and x.stride(0) == ldx and x.dtype == torch.float32
        y = torch.empty((T, d), dtype=torch.float32, device=x.device)
        mean = torch.empty((T,), dtype=torch.float32, device=x.device)
        rstd = torch.empty_like(mean)
        _call("kv_layernorm_fwd", lib.kv_layernorm_fwd, x.data_ptr(), L.ptr(gamma), L.ptr(beta), y.data_ptr(), mean.data_ptr(),
                                     rstd.data_ptr(), T, d, ldx, eps, _stream(), work=8.0 * T * d)
        ctx.save_for_backward(x, gamma, mean, rstd)
        ctx.eps, ctx.ldx = eps, ldx
        ctx.beta = beta          # identity only (gradient sink lookup); beta's values are not needed in backward
        ctx.fork = fork
        if fork:
            ctx.set_materialize_grads(False)
            return y, x.view_as(x)
        return y

    @staticmethod
    def backward(ctx, dy, dskip=None):
        lib = L.load()
        x, gamma, mean, rstd = ctx.saved_tensors
        T, d = x.shape
        if dy is None:                      # only the skip branch was used
            return dskip, None, None, None, None, None
        dy = dy.contiguous()
        if dskip is not None:
            dskip = dskip.contiguous()
            assert dskip.shape == (T, d)
        dx = torch.empty((T, d), dtype=torch.float32, device=x.device)
        # the kernel ADDS its d gamma / d beta partial sums: with a gradient sink they go straight into the
        # parameters' flat gradient views (no zero-fill, no autograd accumulation launch)
        sink = GRAD_SINK
        beta = ctx.beta
        tg = sink.target(gamma) if sink is not None and gamma is not None and gamma.requires_grad else None
        tb = sink.target(beta) if sink is not None and beta is not None and beta.requires_grad else None
        direct = tg is not None and tb is not None and tg.is_contiguous() and tb.is_contiguous()
        if direct:
            dgamma, dbeta = tg, tb
        else:
            dgamma = torch.zeros((d,), dtype=torch.float32, device=x.device) if gamma is not None else None
            dbeta = torch.zeros((d,), dtype=torch.float32, device=x.device) if gamma is not None else None
        _call("kv_layernorm_bwd", lib.kv_layernorm_bwd, x.data_ptr(), dy.data_ptr(), L.ptr(gamma), mean.data_ptr(),
                                     rstd.data_ptr(), L.ptr(dskip), dx.data_ptr(), L.ptr(dgamma), L.ptr(dbeta),
                                     T, d, ctx.ldx, ctx.eps, _stream(), work=(12.0 if dskip is None else 16.0) * T * d)
        if direct:
            sink.done(gamma)
            sink.done(beta)
            return dx, None, None, None, None, None
        return dx, dgamma, dbeta, None, None, None


def layer_norm(x2d, gamma, beta, eps=1e-5):
    """nn.LayerNorm over the last dim of a 2-D (possibly row-strided) fp32 matrix."""
    if x2d.stride(1) != 1:
        x2d = x2d.contiguous()
    return _LayerNormFn.apply(x2d, gamma, beta, eps, x2d.stride(0))


def layer_norm_fork(x2d, gamma, beta, eps=1e-5):
    """(LayerNorm(x), x): use the second value for the skip connection around the normalised branch
    (x + f(LN(x)), model.py:34,36); the backward then fuses the skip gradient into the LayerNorm gradient."""
    if x2d.stride(1) != 1 or x2d.stride(0) != x2d.shape[1]:
        x2d = x2d.contiguous()
    return _LayerNormFn.apply(x2d, gamma, beta, eps, x2d.stride(0), True)


class _PatchifyFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, images, n_patches):
        lib = L.load()
        _require_cuda(images)
        B, Cc, H, W = images.shape
        ph, pw = H // n_patches, W // n_patches
        patches = torch.empty((B, n_patches * n_patches, Cc * ph * pw), dtype=torch.float32,
                              device=images.device)
        _call("kv_patchify", lib.kv_patchify, images.data_ptr(), patches.data_ptr(), B, Cc, H, W, n_patches, _stream(), work=8.0 * B * Cc * H * W)
        ctx.dims = (B, Cc, H, W, n_patches)
        return patches

    @staticmethod
    def backward(ctx, dpatches):
        lib = L.load()
        B, Cc, H, W, n_patches = ctx.dims
        dpatches = dpatches.contiguous()
        dimg = torch.empty((B, Cc, H, W), dtype=torch.float32, device=dpatches.device)
        _call("kv_patchify_bwd", lib.kv_patchify_bwd, dpatches.data_ptr(), dimg.data_ptr(), B, Cc, H, W, n_patches, _stream(), work=8.0 * B * Cc * H * W)
        return dimg, None


def patchify(images, n_patches):
    """VisionTransformer.patchify (model.py:111-126): (B,C,H,W) -> (B, np^2, C*ph*pw) float32."""
    return _PatchifyFn.apply(images.contiguous().float(), n_patches)


_CE_WORKSPACE = {}


def _ce_workspace(device):
    # keyed by (device, stream): the kernel's arrival counter lives in the workspace, so two launches that may
    # overlap (train and eval streams) must not share it
    key = (device, _stream())
    ws = _CE_WORKSPACE.get(key)
    if ws is None:
        ws = torch.zeros((L.load().kv_cross_entropy_workspace_floats(),), dtype=torch.float32, device=device)
        _CE_WORKSPACE[key] = ws
    return ws


def cross_entropy_with_grad(logits, labels, ignore_index=-100, need_grad=True, probs_out=None, pred_out=None,
                            confusion=None):
    """nn.CrossEntropyLoss() (train.py:22,36) and its gradient w.r.t. the logits in ONE pass:
    returns (loss: 0-dim tensor, dlogits or None).  labels: int64 class indices.
    Optional metric outputs written by the same launch (train.py:42-44): ``probs_out`` [B, C] fp32 softmax,
    ``pred_out`` [B] int64 argmax, ``confusion`` [C, C] int32 counts incremented at (label, prediction)."""
    lib = L.load()
    _require_cuda(logits, labels)
    assert logits.dim() == 2 and logits.dtype == torch.float32
    assert labels.shape == (logits.shape[0],) and labels.dtype == torch.int64
    logits, labels = logits.contiguous(), labels.contiguous()
    B, Cn = logits.shape
    _require_cuda(probs_out, pred_out, confusion)
    assert probs_out is None or (probs_out.shape == (B, Cn) and probs_out.dtype == torch.float32 and probs_out.is_contiguous())
    assert pred_out is None or (pred_out.shape == (B,) and pred_out.dtype == torch.int64 and pred_out.is_contiguous())
    assert confusion is None or (confusion.shape == (Cn, Cn) and confusion.dtype == torch.int32 and confusion.is_contiguous())
    loss = torch.empty((), dtype=torch.float32, device=logits.device)
    dz = torch.empty_like(logits) if need_grad else None
    _call("kv_cross_entropy", lib.kv_cross_entropy, logits.data_ptr(), labels.data_ptr(), B, Cn, int(ignore_index),
          loss.data_ptr(), L.ptr(dz), _ce_workspace(logits.device).data_ptr(), L.ptr(probs_out), L.ptr(pred_out),
          L.ptr(confusion), _stream(), work=8.0 * B * Cn)
    return loss, dz


class _CrossEntropyFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, labels, ignore_index):
        loss, dz = cross_entropy_with_grad(logits, labels, ignore_index, need_grad=logits.requires_grad)
        ctx.save_for_backward(dz)
        return loss

    @staticmethod
    def backward(ctx, dloss):
        (dz,) = ctx.saved_tensors
        return dz * dloss, None, None


def cross_entropy(logits, labels, ignore_index=-100):
    """Drop-in for ``nn.CrossEntropyLoss()(logits, labels)`` on CUDA tensors (mean reduction)."""
    return _CrossEntropyFn.apply(logits, labels, ignore_index)


def adam_step(param, grad, exp_avg, exp_avg_sq, state, lr, beta1=0.9, beta2=0.999, eps=1e-8, grad_scale=1.0,
              zero_grad=False):
    """One torch.optim.Adam update (train.py:23,40) of flat fp32 buffers, in place; ``state`` = 3 device doubles
    (step count and bias corrections, advanced on the device)."""
    lib = L.load()
    _require_cuda(param, grad, exp_avg, exp_avg_sq, state)
    n = param.numel()
    assert grad.numel() == n and exp_avg.numel() == n and exp_avg_sq.numel() == n
    assert state.dtype == torch.float64 and state.numel() >= 3
    for t in (param, grad, exp_avg, exp_avg_sq):
        assert t.dtype == torch.float32 and t.is_contiguous()
    _call("kv_adam_step", lib.kv_adam_step, param.data_ptr(), grad.data_ptr(), exp_avg.data_ptr(), exp_avg_sq.data_ptr(),
          n, lr, beta1, beta2, eps, grad_scale, 1 if zero_grad else 0, state.data_ptr(), _stream(), work=32.0 * n)


class _AttnGeneralFn(torch.autograd.Function):
    """FlashAttentionFunction (utils.py:137-295): q [B*Nq, >=H*dh], k/v [B*Nk, ...] column views allowed."""

    @staticmethod
    def forward(ctx, q, k, v, mask, causal, B, Nq, Nk, H, dh):
        lib = L.load()
        _require_cuda(q, k, v, mask)
        for t in (q, k, v):
            assert t.dim() == 2 and t.dtype == torch.float32 and t.stride(1) == 1
        out = torch.empty((B * Nq, H * dh), dtype=torch.float32, device=q.device)
        lse = torch.empty((B, H, Nq), dtype=torch.float32, device=q.device)
        a = L.AttnGeneral(B, H, Nq, Nk, dh, q.stride(0), k.stride(0), v.stride(0), H * dh, 1 if causal else 0, 0, 0, 0)
        if mask is not None:
            assert mask.dtype == torch.bool and mask.dim() == 4 and mask.shape[-1] == Nk and mask.stride(3) == 1
            a.mask_stride_b, a.mask_stride_h, a.mask_stride_q = mask.stride(0), mask.stride(1), mask.stride(2)
        _call("kv_attn_general_fwd", lib.kv_attn_general_fwd, C.byref(a), q.data_ptr(), k.data_ptr(), v.data_ptr(),
              L.ptr(mask), out.data_ptr(), lse.data_ptr(), _stream(), work=4.0 * B * H * Nq * Nk * dh)
        ctx.a = a
        ctx.save_for_backward(q, k, v, mask, out, lse)
        return out

    @staticmethod
    def backward(ctx, dout):
        lib = L.load()
        q, k, v, mask, out, lse = ctx.saved_tensors
        a = ctx.a
        dout = dout.contiguous()
        # gradients in the layouts of q / k / v (k and v are usually the halves of one to_kv output)
        dq = torch.empty_strided(q.shape, q.stride(), dtype=torch.float32, device=q.device)
        if k.stride(0) == v.stride(0) and v.data_ptr() - k.data_ptr() == 4 * k.shape[1] and k.stride(0) == 2 * k.shape[1]:
            dkv = torch.empty((k.shape[0], 2 * k.shape[1]), dtype=torch.float32, device=q.device)
            dk, dv = dkv[:, :k.shape[1]], dkv[:, k.shape[1]:]
        else:
            dk = torch.empty_strided(k.shape, k.stride(), dtype=torch.float32, device=q.device)
            dv = torch.empty_strided(v.shape, v.stride(), dtype=torch.float32, device=q.device)
        _call("kv_attn_general_bwd", lib.kv_attn_general_bwd, C.byref(a), q.data_ptr(), k.data_ptr(), v.data_ptr(),
              L.ptr(mask), out.data_ptr(), dout.data_ptr(), lse.data_ptr(), dq.data_ptr(), dk.data_ptr(),
              dv.data_ptr(), _stream(), work=10.0 * a.B * a.H * a.Nq * a.Nk * a.d_head)
        return dq, dk, dv, None, None, None, None, None, None, None


def attention_general(q, k, v, B, Nq, Nk, H, dh, mask=None, causal=False):
    """Attention of the FlashAttention module: q [B*Nq, H*dh], k / v [B*Nk, H*dh] (column views of a wider
    matrix are fine), mask None or a bool tensor broadcastable to [B, H, Nq, Nk] -> out [B*Nq, H*dh]."""
    if mask is not None:
        if mask.dim() == 2:                       # 'b n' key mask (utils.py:154-155)
            mask = mask[:, None, None, :]
        mask = mask.to(torch.bool)
        if mask.shape[-1] != Nk or mask.stride(-1) != 1:
            mask = mask.expand(*mask.shape[:-1], Nk).contiguous()
        mask = mask.expand(B, H, Nq, Nk)
    return _AttnGeneralFn.apply(q, k, v, mask, bool(causal), B, Nq, Nk, H, dh)


def image_augment(images_u8_hwc, mean, std, pad=0, flip=None, top=None, left=None):
    """The transform of train.py:99-109 on the GPU: uint8 [B, H, W, C] -> fp32 [B, C, H, W]
    (horizontal flip where ``flip`` [B] bool/uint8 is set, zero padding ``pad`` and a crop at offsets
    ``top`` / ``left`` [B] int32 in [0, 2*pad], /255, (v - mean) / std).  flip/top/left None = the test transform."""
    lib = L.load()
    _require_cuda(images_u8_hwc, mean, std, flip, top, left)
    assert images_u8_hwc.dtype == torch.uint8 and images_u8_hwc.dim() == 4 and images_u8_hwc.is_contiguous()
    B, H, W, Cc = images_u8_hwc.shape
    assert mean.numel() == Cc and std.numel() == Cc and mean.dtype == torch.float32 and std.dtype == torch.float32
    if flip is not None:
        flip = flip.to(torch.uint8).contiguous()
    for t in (top, left):
        assert t is None or (t.dtype == torch.int32 and t.shape == (B,) and t.is_contiguous())
    out = torch.empty((B, Cc, H, W), dtype=torch.float32, device=images_u8_hwc.device)
    _call("kv_image_augment", lib.kv_image_augment, images_u8_hwc.data_ptr(), L.ptr(flip), L.ptr(top), L.ptr(left),
          B, H, W, Cc, pad, mean.data_ptr(), std.data_ptr(), out.data_ptr(), _stream(), work=5.0 * B * Cc * H * W)
    return out
