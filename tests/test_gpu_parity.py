"""GPU parity: the CUDA path (through the C ABI) against the golden vectors generated from the
unmodified reference, and against the oracle on larger seeded inputs.

Tolerance (BASELINE.json north_star): fp32 path within 1e-5 relative, measured norm-wise
(max|a-b| / max|b| per tensor); layout pieces bit-exact.
"""
import numpy as np
import pytest
import torch

from conftest import load_golden, relerr

pytestmark = pytest.mark.gpu

TOL_FP32 = 1e-5
TOL_GRAD = 2e-5  # gradients pass through one more fp32 reduction (over T) than outputs
ATOL_GRAD = 1e-8  # floor for gradients that are a cancelling sum (SineKAN d freq ~1e-5 from terms ~1e-2)


def grad_close(got, ref, tol=TOL_GRAD):
    got, ref = got.detach().double().cpu(), ref.detach().double().cpu()
    return (got - ref).abs().max().item() <= tol * ref.abs().max().item() + ATOL_GRAD


def _dev():
    return torch.device("cuda:0")


def _build(case):
    import kan_vit_b200 as K
    mk = {
        "layer_effkan": lambda: K.KANLinear(16, 24),
        "layer_effkan_g10k2_nonuniform": lambda: K.KANLinear(8, 12, grid_size=10, spline_order=2,
                                                             enable_standalone_scale_spline=False),
        "layer_cheby": lambda: K.ChebyKANLayer(16, 24, 4),
        "layer_cheby_deg7": lambda: K.ChebyKANLayer(8, 8, 7),
        "layer_fastkan": lambda: K.FastKANLayer(16, 24),
        "layer_sine": lambda: K.SineKANLayer(16, 24),
        "layer_sine_g4": lambda: K.SineKANLayer(8, 8, grid_size=4),
        "layer_fourier": lambda: K.NaiveFourierKANLayer(16, 24, 5),
        "block_efficientkan": lambda: K.TransformerBlock(16, 2, feedforward_dim=64, attn_type="efficientkan"),
        "block_efficientkan_kanmlp": lambda: K.TransformerBlock(16, 2, feedforward_dim=64,
                                                                attn_type="efficientkan", kan_mlp=True),
    }
    if case.startswith("msa_"):
        return K.MSA(16, n_heads=2, type=case[4:])
    if case.startswith("vit_"):
        return K.VisionTransformer((1, 8, 8), n_patches=2, n_blocks=2, d_hidden=16, n_heads=2, out_d=10,
                                   type=case[4:])
    return mk[case]()


CASES = ["layer_effkan", "layer_effkan_g10k2_nonuniform", "layer_cheby", "layer_cheby_deg7",
         "layer_fastkan", "layer_sine", "layer_sine_g4", "layer_fourier",
         "msa_vanilla", "msa_efficientkan", "msa_cheby", "msa_fast", "msa_sine",
         "block_efficientkan", "block_efficientkan_kanmlp",
         "vit_vanilla", "vit_efficientkan", "vit_fast", "vit_sine", "vit_cheby", "vit_flash-attn"]


@pytest.mark.parametrize("case", CASES)
def test_matches_reference_golden(case):
    g = load_golden(case)
    m = _build(case)
    missing, unexpected = m.load_state_dict(g.sd(), strict=False)
    assert not unexpected, unexpected
    assert all(k.endswith("pos_embeddings") for k in missing), missing
    m = m.to(_dev())
    x = g.t("x", _dev()).requires_grad_(True)
    y = m(x)
    ref_y = g.t("y")
    assert y.shape == ref_y.shape
    assert relerr(y, ref_y) < TOL_FP32
    y.backward(g.t("dy", _dev()))
    assert relerr(x.grad, g.t("gx")) < TOL_GRAD
    params = dict(m.named_parameters())
    # gradients that are a cancelling sum (q/k projections under a nearly uniform softmax) carry the fp32 rounding
    # of the large terms they cancel: allow 1e-7 of the model's largest gradient as an absolute floor
    floor = 1e-7 * max(ref.abs().max().item() for ref in g.grads().values())
    for k, ref in g.grads().items():
        got = params[k].grad
        assert got is not None, k
        if ref.abs().max() < 1e-6:  # analytically zero (key bias under softmax)
            assert got.abs().max() < 1e-5, k
            continue
        assert grad_close(got, ref) or (got.cpu() - ref).abs().max().item() <= floor, (k, relerr(got, ref))


def test_patchify_and_pos_bit_exact():
    import kan_vit_b200 as K
    z = np.load(__import__("os").path.join(__import__("conftest").GOLDEN, "layout.npz"))
    v = K.VisionTransformer((3, 32, 32), n_patches=4, n_blocks=1, d_hidden=24, n_heads=2).to(_dev())
    got = v.patchify(torch.from_numpy(z["img_3x32"]).to(_dev()), 4).cpu()
    assert torch.equal(got, torch.from_numpy(z["patches_3x32_np4"]))
    assert torch.equal(v.pos_embeddings.cpu(), torch.from_numpy(z["pos_17x24"]))
    v7 = K.VisionTransformer((1, 28, 28), n_patches=7, n_blocks=1, d_hidden=64, n_heads=2).to(_dev())
    got = v7.patchify(torch.from_numpy(z["img_1x28"]).to(_dev()), 7).cpu()
    assert torch.equal(got, torch.from_numpy(z["patches_1x28_np7"]))
    assert torch.equal(v7.pos_embeddings.cpu(), torch.from_numpy(z["pos_50x64"]))


# ---------------------------------------------------------------------------------------------
# larger seeded inputs against the oracle (CPU, fp32), incl. ragged T and edge values
# ---------------------------------------------------------------------------------------------
def _oracle_layer(kind, m, x):
    from oracle import kan_oracle as O
    sd = {k: v.detach().cpu().clone().requires_grad_(v.is_floating_point()) for k, v in m.state_dict().items()}
    xc = x.detach().cpu().clone().requires_grad_(True)
    y = O.layer_forward(kind, xc, sd, "")
    return y, xc, sd


LAYER_SWEEP = [
    ("efficientkan", lambda K: K.KANLinear(37, 70), 37),
    ("efficientkan", lambda K: K.KANLinear(64, 64, grid_size=20), 64),
    ("cheby", lambda K: K.ChebyKANLayer(33, 65, 6), 33),
    ("fast", lambda K: K.FastKANLayer(48, 40), 48),
    ("sine", lambda K: K.SineKANLayer(40, 72, grid_size=10), 40),
    ("fourier", lambda K: K.NaiveFourierKANLayer(24, 40, 7), 24),
    ("vanilla", lambda K: torch.nn.Linear(50, 77), 50),
]


@pytest.mark.parametrize("kind,make,n_in", LAYER_SWEEP)
@pytest.mark.parametrize("T", [1, 63, 1000])
def test_layer_vs_oracle(kind, make, n_in, T):
    import kan_vit_b200 as K
    from kan_vit_b200 import ops
    from kan_vit_b200._dispatch import make_spec, slots_of
    torch.manual_seed(7)
    m = make(K).to(_dev())
    x = (torch.randn(T, n_in, generator=torch.Generator().manual_seed(T)) * 1.2)
    if kind == "efficientkan" and x.numel() >= 6:  # knots, span edges, outside the span
        x.view(-1)[:6] = torch.tensor([-2.2, 2.2, -1.0, 0.2, 3.0, -2.6])
    xg = x.to(_dev()).requires_grad_(True)
    if kind == "vanilla":
        y = ops.kan_apply(xg, make_spec([m]), [slots_of(m)])
    else:
        y = m(xg)
    yo, xc, sd = _oracle_layer(kind, m, x)
    assert y.shape == yo.shape
    assert relerr(y, yo) < TOL_FP32
    dy = torch.randn(y.shape, generator=torch.Generator().manual_seed(99))
    y.backward(dy.to(_dev()))
    yo.backward(dy)
    assert relerr(xg.grad, xc.grad) < TOL_GRAD
    for k, p in m.named_parameters():
        if not p.requires_grad:
            continue
        assert grad_close(p.grad, sd[k].grad), (k, relerr(p.grad, sd[k].grad))


def test_empty_batch():
    import kan_vit_b200 as K
    m = K.KANLinear(8, 8).to(_dev())
    y = m(torch.empty(0, 8, device=_dev()))
    assert y.shape == (0, 8)


def test_shape_mismatch_raises_like_reference():
    import kan_vit_b200 as K
    m = K.KANLinear(8, 8).to(_dev())
    with pytest.raises(AssertionError):
        m(torch.zeros(4, 9, device=_dev()))
    with pytest.raises(ValueError):
        K.VisionTransformer((1, 28, 28), type="nope")
    with pytest.raises(AssertionError):
        K.VisionTransformer((1, 28, 28), n_patches=5)


def _min_relu_margin(v, img, kind):
    """min |pre-activation| over every ReLU unit of the model, from the float64 oracle."""
    from oracle import kan_oracle as O
    seen = []
    orig = O.F.relu

    def spy(x, *a, **k):
        seen.append(x.detach().abs().min().item())
        return orig(x, *a, **k)

    O.F.relu = spy
    try:
        sd = {k: (t.detach().double() if t.is_floating_point() else t) for k, t in v.state_dict().items()}
        with torch.no_grad():
            O.vit_forward(img.double(), sd, n_patches=7, n_blocks=4, n_heads=2, kind=kind)
    finally:
        O.F.relu = orig
    return min(seen) if seen else 1.0


@pytest.mark.parametrize("kind", ["efficientkan", "cheby", "fast", "sine", "vanilla"])
def test_vit_cfg_shapes_vs_oracle(kind):
    """BASELINE config 1/2 geometry (1x28x28, n_patches 7, d 64, depth 4, heads 2) at batch 8.

    Through 4 blocks the fp32 oracle itself drifts from exact arithmetic (summation order, ReLU
    units whose pre-activation is ~0 flip their mask), so gradients are judged against the
    float64 oracle: our error must be within 5e-5, or no worse than twice the fp32 oracle's own
    error against float64 ("who is closer to fp64 settles the dispute")."""
    import kan_vit_b200 as K
    from oracle import kan_oracle as O
    img = torch.randn(8, 1, 28, 28, generator=torch.Generator().manual_seed(5))
    # A ReLU unit whose pre-activation is within fp32 noise of 0 makes the gradient discontinuous
    # (seed 3 / efficientkan has one at 5.9e-9): pick the first seed whose float64 oracle keeps
    # every pre-activation of ff.0 at least 1e-6 away from zero.
    for seed in range(3, 40):
        torch.manual_seed(seed)
        v = K.VisionTransformer((1, 28, 28), 7, 4, 64, 2, 10, kind)
        if _min_relu_margin(v, img, kind) > 1e-6:
            break
    v = v.to(_dev())
    y = v(img.to(_dev()))
    dy = torch.randn(y.shape, generator=torch.Generator().manual_seed(6))
    y.backward(dy.to(_dev()))

    def run_oracle(dtype):
        sd = {k: t.detach().cpu().to(dtype if t.is_floating_point() else t.dtype).clone()
              .requires_grad_(t.is_floating_point()) for k, t in v.state_dict().items()}
        yo = O.vit_forward(img.to(dtype), sd, n_patches=7, n_blocks=4, n_heads=2, kind=kind)
        yo.backward(dy.to(dtype))
        return yo, sd

    y32, sd32 = run_oracle(torch.float32)
    y64, sd64 = run_oracle(torch.float64)
    assert relerr(y, y32) < TOL_FP32 or relerr(y, y64) <= 2 * relerr(y32, y64)
    for k, p in v.named_parameters():
        if not p.requires_grad:
            continue
        truth = sd64[k].grad
        if truth.abs().max() < 1e-6:
            assert p.grad.abs().max() < 1e-5, k
            continue
        ours, theirs = relerr(p.grad, truth), relerr(sd32[k].grad, truth)
        # SineKAN's d freq (G values per layer) is a cancelling sum over all T*in terms, each
        # ~1e3 times the result: fp32 summation order alone moves it by ~1e-4 relative (the fp32
        # oracle is itself ~2e-5..1e-4 off float64 there), so it gets the looser bound.
        tol = 3e-4 if k.endswith(".freq") else 5e-5
        assert ours < tol or ours <= 2 * theirs, (k, ours, theirs)


def test_full_size_row_independence_cfg2():
    """Size-independent property at BASELINE config 2's full batch (cheby ViT, B=1024): every op of
    the path is per-sample, so the logits of a sub-batch must equal the same rows of the full
    batch bit for bit, and weight gradients must add over a split of the batch."""
    import kan_vit_b200 as K
    torch.manual_seed(11)
    v = K.VisionTransformer((1, 28, 28), 7, 4, 64, 2, 10, "cheby").to(_dev())
    img = torch.randn(1024, 1, 28, 28, device=_dev())
    with torch.no_grad():
        full = v(img)
        part = v(img[100:228])
    assert torch.equal(full[100:228], part)
    dy = torch.randn_like(full)
    v.zero_grad()
    v(img).backward(dy)
    g_full = {k: p.grad.clone() for k, p in v.named_parameters() if p.grad is not None}
    v.zero_grad()
    v(img[:512]).backward(dy[:512])
    v(img[512:]).backward(dy[512:])
    for k, p in v.named_parameters():
        if p.grad is None or g_full[k].abs().max() < 1e-6:
            continue
        assert relerr(p.grad, g_full[k]) < 1e-4, k


# ---- FlashAttention module / type="flash-attn"  ----

FLASH = ["flash_self", "flash_keymask", "flash_causal", "flash_cross_causal", "flash_cross_mask",
         "flash_fullmask_causal_buckets"]


@pytest.mark.parametrize("case", FLASH)
def test_flash_attention_matches_reference_golden(case):
    import kan_vit_b200 as K
    g = load_golden(case)
    m = K.FlashAttention(dim=16, heads=2, dim_head=8, causal=bool(g.raw["causal"]))
    m.load_state_dict(g.sd())
    m = m.to(_dev())
    x = g.t("x", _dev()).requires_grad_(True)
    ctx = g.t("context", _dev()) if "context" in g.raw else None
    mask = g.t("mask", _dev()) if "mask" in g.raw else None
    y = m(x, context=ctx, mask=mask, q_bucket_size=3, k_bucket_size=2)
    assert relerr(y, g.t("y")) < TOL_FP32
    y.backward(g.t("dy", _dev()))
    assert relerr(x.grad, g.t("gx")) < TOL_GRAD
    params = dict(m.named_parameters())
    for k, ref in g.grads().items():
        assert grad_close(params[k].grad, ref), (k, relerr(params[k].grad, ref))


@pytest.mark.parametrize("Nq,Nk,H,dh,causal,masked", [(197, 197, 3, 64, False, False), (197, 197, 2, 64, True, True),
                                                      (50, 120, 4, 32, True, False), (130, 40, 2, 24, False, True),
                                                      (1, 1, 1, 8, True, False)])
def test_flash_attention_vs_oracle(Nq, Nk, H, dh, causal, masked):
    import kan_vit_b200 as K
    from oracle import kan_oracle as O
    torch.manual_seed(Nq + Nk)
    B, dim = 3, 48
    m = K.FlashAttention(dim=dim, heads=H, dim_head=dh, causal=causal)
    x = torch.randn(B, Nq, dim)
    ctx = torch.randn(B, Nk, dim) if Nk != Nq else None
    mask = None
    if masked:
        mask = torch.rand(B, Nk) < 0.8
        mask[0, : Nk // 2] = False
    dy = torch.randn(B, Nq, dim)
    sd = {k: v.detach().clone().requires_grad_(True) for k, v in m.state_dict().items()}
    xo = x.clone().requires_grad_(True)
    yo = O.flash_attention_forward(xo, sd, "", H, context=ctx, mask=mask, causal=causal)
    yo.backward(dy)
    m = m.to(_dev())
    xg = x.to(_dev()).requires_grad_(True)
    y = m(xg, context=None if ctx is None else ctx.to(_dev()), mask=None if mask is None else mask.to(_dev()))
    assert relerr(y, yo) < TOL_FP32
    y.backward(dy.to(_dev()))
    assert relerr(xg.grad, xo.grad) < TOL_GRAD
    for k, p in m.named_parameters():
        if sd[k].grad.abs().max() < 1e-6:      # analytically zero (a single key: softmax = 1, no q / k gradient)
            assert p.grad.abs().max().item() < 1e-5, k
            continue
        assert grad_close(p.grad, sd[k].grad), (k, relerr(p.grad, sd[k].grad))


def test_flash_attention_api_like_reference():
    import kan_vit_b200 as K
    with pytest.raises(TypeError):
        K.FlashAttention(16)                      # keyword-only, attention.py:14-16
    m = K.FlashAttention(dim=16)
    assert m.to_q.weight.shape == (512, 16) and m.to_kv.weight.shape == (1024, 16) and m.to_out.weight.shape == (16, 512)
    assert m.to_q.bias is None and m.to_out.bias is None
    v = K.VisionTransformer((1, 8, 8), n_patches=2, n_blocks=2, d_hidden=16, n_heads=2, out_d=10, type="flash-attn")
    assert sorted(k for k in v.state_dict() if k.startswith("blocks.0")) == [
        "blocks.0.to_kv.weight", "blocks.0.to_out.weight", "blocks.0.to_q.weight"]
    v = v.to(_dev())
    with K.precision("bf16"):
        y = v(torch.randn(4, 1, 8, 8, device=_dev()))
    assert y.shape == (4, 10) and torch.isfinite(y).all()


@pytest.mark.parametrize("mode,tol", [("fp32", 2e-4), ("bf16", 2e-2)])
def test_update_grid_on_device_then_forward(mode, tol):
    """the maintenance ops run on the device (torch sort / lstsq on CUDA tensors) and the
    kernels honour the adapted, per-feature non-uniform grid: outputs against the reference's after update_grid.
    (The refit is a least-squares solve, so the fp32 tolerance is the solve's, not the kernel's 1e-5; the kernel
    itself is compared at 1e-5 with the reference's own refit weights loaded.)"""
    import kan_vit_b200 as K
    g = load_golden("maint_effkan")
    m = K.KANLinear(6, 5)
    m.load_state_dict(g.sd())
    m = m.to(_dev())
    with K.precision(mode):
        assert relerr(m(g.t("probe", _dev())), g.t("y_before")) < (TOL_FP32 if mode == "fp32" else tol)
        m.update_grid(g.t("x", _dev()))
        assert relerr(m.grid, g.t("grid_after")) < 1e-6
        # the refit reproduces the layer on the samples it was fitted to (update_grid's contract)
        assert relerr(m(g.t("x", _dev())), g.t("y_after_x")) < tol
        assert relerr(m(g.t("probe", _dev())), g.t("y_after")) < 10 * tol
        with torch.no_grad():
            m.spline_weight.copy_(g.t("spline_weight_after", _dev()))
        assert relerr(m(g.t("probe", _dev())), g.t("y_after")) < (TOL_FP32 if mode == "fp32" else tol)


@pytest.mark.parametrize("T,d", [(5, 16), (1000, 64), (333, 192), (2049, 768), (70, 1024), (64, 520), (33, 1028), (40, 770),
                                 (17, 4)])
def test_layer_norm_vs_torch(T, d):
    """nn.LayerNorm (model.py:22,24,107) forward/backward, register-resident kernels (d % 4 == 0, d <= 1024) and
    the generic ones, contiguous and row-strided inputs."""
    from kan_vit_b200 import ops
    dev = _dev()
    g = torch.Generator().manual_seed(T * d)
    w = (1.0 + 0.1 * torch.randn(d, generator=g)).to(dev).requires_grad_(True)
    b = (0.1 * torch.randn(d, generator=g)).to(dev).requires_grad_(True)
    for strided in (False, True):
        base = (2.0 * torch.randn(T, 2 * d if strided else d, generator=g) + 0.5).to(dev).requires_grad_(True)
        x = base[:, :d] if strided else base
        dy = torch.randn(T, d, generator=g).to(dev)
        y = ops.layer_norm(x, w, b)
        gx, gw, gb = torch.autograd.grad(y, (base, w, b), dy)
        yr = torch.nn.functional.layer_norm(x.double(), (d,), w.double(), b.double(), 1e-5)
        rx, rw, rb = torch.autograd.grad(yr, (base, w, b), dy.double())
        assert relerr(y, yr) < TOL_FP32
        assert relerr(gx, rx) < TOL_GRAD and relerr(gw, rw) < TOL_GRAD and relerr(gb, rb) < TOL_GRAD


@pytest.mark.parametrize("T,d", [(300, 768), (37, 100), (5, 16)])
def test_layer_norm_fork_fuses_skip_gradient(T, d):
    """(LN(x), x) with both branches used: one backward call receives both gradients and the kernel adds the skip
    gradient into dX (model.py:34,36 residual structure); also each branch alone."""
    from kan_vit_b200 import ops
    dev = _dev()
    g = torch.Generator().manual_seed(d)
    w = (1.0 + 0.1 * torch.randn(d, generator=g)).to(dev).requires_grad_(True)
    b = (0.1 * torch.randn(d, generator=g)).to(dev).requires_grad_(True)
    x = torch.randn(T, d, generator=g).to(dev).requires_grad_(True)
    a1, a2 = torch.randn(T, d, generator=g).to(dev), torch.randn(T, d, generator=g).to(dev)
    for use_y, use_s in ((True, True), (True, False), (False, True)):
        y, s = ops.layer_norm_fork(x, w, b)
        assert torch.equal(s, x)
        out = 0.0
        if use_y:
            out = out + (y * a1).sum()
        if use_s:
            out = out + (s * a2).sum()
        got = torch.autograd.grad(out, (x, w, b), allow_unused=True)
        yr = torch.nn.functional.layer_norm(x.double(), (d,), w.double(), b.double(), 1e-5)
        ref_out = ((yr * a1).sum() if use_y else 0.0) + ((x.double() * a2).sum() if use_s else 0.0)
        ref = torch.autograd.grad(ref_out, (x, w, b), allow_unused=True)
        for a, r in zip(got, ref):
            if r is None:
                assert a is None or a.abs().max().item() == 0.0
            else:
                assert relerr(a, r) < TOL_GRAD
