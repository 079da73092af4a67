"""Pins oracle/kan_oracle.py to the reference's own outputs (tests/golden, made by
oracle/make_golden.py from the unmodified /root/reference).  CPU only."""
import pytest
import torch

from conftest import load_golden, relerr
from oracle import kan_oracle as O

TOL = 2e-6  # same fp32 arithmetic, restated; only einsum/linear summation order may differ


def _leafify(sd):
    return {k: (v.clone().requires_grad_(True) if v.is_floating_point() else v) for k, v in sd.items()}


def _check(g, y, x, sd, tol=TOL, skip_nograd=()):
    assert y.shape == g.t("y").shape
    assert relerr(y, g.t("y")) < tol
    y.backward(g.t("dy"))
    assert relerr(x.grad, g.t("gx")) < 10 * tol
    for k, ref in g.grads().items():
        assert sd[k].grad is not None, k
        if ref.abs().max() < 1e-6:  # analytically zero (e.g. key bias under softmax): noise only
            assert sd[k].grad.abs().max() < 1e-6, k
            continue
        assert relerr(sd[k].grad, ref) < 10 * tol, k


LAYERS = [
    ("layer_effkan", "efficientkan"), ("layer_cheby", "cheby"), ("layer_cheby_deg7", "cheby"),
    ("layer_fastkan", "fast"), ("layer_sine", "sine"), ("layer_sine_g4", "sine"),
    ("layer_fourier", "fourier"),
]


@pytest.mark.parametrize("case,kind", LAYERS)
def test_layer_matches_reference(case, kind):
    g = load_golden(case)
    sd = _leafify(g.sd())
    x = g.t("x").requires_grad_(True)
    y = O.layer_forward(kind, x, sd, "")
    _check(g, y, x, sd)
    # and the fp64 ground truth: fp32 oracle is as close to it as the reference is
    assert relerr(y, g.t("y64")) < 1e-5


def test_effkan_nonuniform_grid_any_order():
    g = load_golden("layer_effkan_g10k2_nonuniform")
    sd = _leafify(g.sd())
    x = g.t("x").requires_grad_(True)
    y = O.effkan_forward(x, sd["base_weight"], sd["spline_weight"], None, sd["grid"], spline_order=2)
    _check(g, y, x, sd)


def test_bspline_partition_of_unity_and_support():
    grid = O.effkan_default_grid(3)
    x = torch.linspace(-1, 1, 41)[:, None].repeat(1, 3)
    b = O.b_splines(x, grid, 3)
    assert torch.allclose(b.sum(-1), torch.ones(41, 3), atol=1e-6)
    out = O.b_splines(torch.tensor([[2.2, -2.3, 5.0]]), grid, 3)
    assert out.abs().max() == 0  # half-open span, zero outside (effkan.py:115)


def test_default_buffers_match_reference():
    g = load_golden("layer_effkan")
    assert torch.equal(O.effkan_default_grid(16), g.t("sd/grid"))
    s = load_golden("layer_sine")
    assert torch.equal(O.sinekan_phase(16, 5), s.t("sd/phase"))
    assert torch.equal(O.sinekan_default_freq(5), s.t("sd/freq"))
    s4 = load_golden("layer_sine_g4")
    assert torch.equal(O.sinekan_phase(8, 4), s4.t("sd/phase"))


@pytest.mark.parametrize("kind", ["vanilla", "efficientkan", "cheby", "fast", "sine"])
def test_msa_matches_reference(kind):
    g = load_golden(f"msa_{kind}")
    sd = _leafify(g.sd())
    x = g.t("x").requires_grad_(True)
    y = O.msa_forward(x, sd, "", 2, kind)
    _check(g, y, x, sd)


@pytest.mark.parametrize("case,kan_mlp", [("block_efficientkan", False), ("block_efficientkan_kanmlp", True)])
def test_block_matches_reference(case, kan_mlp):
    g = load_golden(case)
    sd = _leafify(g.sd())
    x = g.t("x").requires_grad_(True)
    y = O.block_forward(x, sd, "", 2, "efficientkan", kan_mlp)
    _check(g, y, x, sd, tol=5e-6)


@pytest.mark.parametrize("kind", ["vanilla", "efficientkan", "fast", "sine", "cheby"])
def test_vit_matches_reference(kind):
    g = load_golden(f"vit_{kind}")
    sd = _leafify(g.sd())
    x = g.t("x").requires_grad_(True)
    y = O.vit_forward(x, sd, n_patches=2, n_blocks=2, n_heads=2, kind=kind)
    _check(g, y, x, sd, tol=1e-5)
    assert torch.equal(O.positional_embeddings(5, 16), g.t("pos_embeddings"))


def test_layout_bit_exact():
    g = load_golden("layout")
    assert torch.equal(O.patchify(g.t("img_3x32"), 4), g.t("patches_3x32_np4"))
    assert torch.equal(O.patchify(g.t("img_1x28"), 7), g.t("patches_1x28_np7"))
    assert torch.equal(O.positional_embeddings(17, 24), g.t("pos_17x24"))
    assert torch.equal(O.positional_embeddings(50, 64), g.t("pos_50x64"))


def test_cross_entropy_restatement_matches_torch():
    """train.py:22 uses torch's nn.CrossEntropyLoss: the numpy restatement is pinned against it."""
    import numpy as np
    g = torch.Generator().manual_seed(3)
    for B, Cn, ignore in ((8, 10, False), (37, 100, True), (256, 1000, True)):
        z = (3.0 * torch.randn(B, Cn, generator=g)).requires_grad_(True)
        y = torch.randint(0, Cn, (B,), generator=g)
        if ignore:
            y[::5] = -100
        loss = torch.nn.CrossEntropyLoss()(z, y)
        loss.backward()
        lo, dz = O.cross_entropy_np(z.detach().numpy(), y.numpy())
        assert abs(float(lo) - float(loss)) <= 2e-6 * abs(float(loss))
        assert np.abs(dz - z.grad.numpy()).max() <= 1e-7


def test_adam_restatement_matches_torch():
    """train.py:23 uses torch.optim.Adam(lr): five steps of the numpy restatement follow it."""
    import numpy as np
    g = torch.Generator().manual_seed(4)
    p = torch.randn(1000, generator=g).requires_grad_(True)
    opt = torch.optim.Adam([p], lr=1e-2)
    pn = p.detach().numpy().copy()
    m = np.zeros_like(pn)
    v = np.zeros_like(pn)
    for t in range(1, 6):
        grad = torch.randn(1000, generator=g) * (10.0 ** float(torch.randint(-3, 2, (1,), generator=g)))
        p.grad = grad.clone()
        opt.step()
        pn, m, v = O.adam_step_np(pn, grad.numpy(), m, v, t, lr=1e-2)
        assert np.abs(pn - p.detach().numpy()).max() <= 2e-6 * np.abs(pn).max()


FLASH = ["flash_self", "flash_keymask", "flash_causal", "flash_cross_causal", "flash_cross_mask",
         "flash_fullmask_causal_buckets"]


@pytest.mark.parametrize("case", FLASH)
def test_flash_attention_matches_reference(case):
    """FlashAttention module (attention.py:13-109) incl. its custom backward (utils.py:229-295): key mask with a
    fully hidden sample, causal, more keys than queries (uniform rows), full mask with non-default buckets."""
    g = load_golden(case)
    sd = _leafify(g.sd())
    x = g.t("x").requires_grad_(True)
    ctx = g.t("context") if "context" in g.raw else None
    mask = g.t("mask") if "mask" in g.raw else None
    y = O.flash_attention_forward(x, sd, "", 2, context=ctx, mask=mask, causal=bool(g.raw["causal"]))
    _check(g, y, x, sd, tol=5e-6)


def test_vit_flash_attn_matches_reference():
    g = load_golden("vit_flash-attn")
    sd = _leafify(g.sd())
    x = g.t("x").requires_grad_(True)
    y = O.vit_forward(x, sd, n_patches=2, n_blocks=2, n_heads=2, kind="flash-attn")
    _check(g, y, x, sd, tol=1e-5)


def test_init_state_has_the_reference_keys_and_shapes():
    """oracle.init_state (the initial weights of bench.py's CPU arm, drawn without importing the product package)
    names and shapes every tensor like the reference's state_dict (tests/golden/state_dict_keys.json, recorded from
    the unmodified reference)."""
    import json
    import os
    from conftest import GOLDEN
    with open(os.path.join(GOLDEN, "state_dict_keys.json")) as f:
        keys = json.load(f)
    cfg = {"cfg1_efficientkan": ((1, 28, 28), 7, 4, 64, 2, 10, "efficientkan"),
           "cfg3_fast": ((3, 32, 32), 4, 6, 192, 8, 10, "fast"), "cfg3_sine": ((3, 32, 32), 4, 6, 192, 8, 10, "sine"),
           "cfg2_cheby": ((1, 28, 28), 7, 4, 64, 2, 10, "cheby"), "vanilla": ((1, 28, 28), 7, 2, 64, 2, 10, "vanilla")}
    for name, a in cfg.items():
        got = {k: list(v.shape) for k, v in O.init_state(*a).items()}
        assert got == keys[name], name
