"""Sequences longer than one CTA's shared memory (csrc/attn_tiled.cu): kv_attn_{fwd,bwd} and kv_attn_general_{fwd,bwd}
fall back to the tiled exact fp32 kernels instead of returning KV_ERR_BAD_SHAPE (the reference accepts any length:
attention.py:199-200 materialises the scores, utils.py:137-295 walks buckets).  384x384 / 16 gives N = 577.
Checked through the C ABI against torch in float64 on the same inputs and, at module level, against the oracle.
Tolerances: the fp32 path's 1e-5 (outputs) / 2e-5 (gradients)."""
import ctypes as C

import pytest
import torch

from conftest import relerr

pytestmark = pytest.mark.gpu
TOL, TOL_GRAD = 1e-5, 2e-5


@pytest.mark.parametrize("B,N,H,dh", [(2, 577, 2, 64), (1, 300, 3, 128), (2, 1025, 1, 24), (1, 230, 2, 64)])
def test_msa_attention_any_length_through_the_c_abi(B, N, H, dh):
    """(1, 230, 2, 64): the forward still fits one CTA, the backward does not -- the two routes must agree on lse."""
    from kan_vit_b200 import _lib as L
    lib = L.load()
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(N)
    qkv = torch.randn(B * N, H, 3, dh, generator=g).to(dev)
    res = torch.randn(B * N, H * dh, generator=g).to(dev)
    dout = torch.randn(B * N, H * dh, generator=g).to(dev)
    out = torch.empty(B * N, H * dh, device=dev)
    lse = torch.empty(B, H, N, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    L.check(lib.kv_attn_fwd(qkv.data_ptr(), res.data_ptr(), out.data_ptr(), lse.data_ptr(), B, N, H, dh, st), "kv_attn_fwd")
    dqkv = torch.full_like(qkv, float("nan"))
    L.check(lib.kv_attn_bwd(qkv.data_ptr(), dout.data_ptr(), lse.data_ptr(), dqkv.data_ptr(), B, N, H, dh, st), "kv_attn_bwd")
    torch.cuda.synchronize()
    x = qkv.double().view(B, N, H, 3, dh).permute(3, 0, 2, 1, 4).clone().requires_grad_(True)   # [3][B][H][N][dh]
    s = torch.einsum("bhid,bhjd->bhij", x[0], x[1]) / dh ** 0.5
    o = torch.einsum("bhij,bhjd->bhid", torch.softmax(s, -1), x[2]).permute(0, 2, 1, 3).reshape(B * N, H * dh)
    assert relerr(out - res, o) < TOL
    assert relerr(lse, torch.logsumexp(s, -1)) < TOL
    o.backward(dout.double())
    ref = x.grad.permute(1, 3, 2, 0, 4).reshape(B * N, H, 3, dh)
    assert torch.isfinite(dqkv).all()
    for part, name in enumerate("qkv"):
        assert relerr(dqkv[:, :, part], ref[:, :, part]) < TOL_GRAD, name


def test_msa_module_long_sequence_matches_oracle():
    import kan_vit_b200 as K
    from oracle import kan_oracle as O
    K.set_precision("fp32")
    torch.manual_seed(3)
    B, N, d, H = 2, 577, 64, 2
    m = K.MSA(d, n_heads=H, type="vanilla")
    x = torch.randn(B, N, d)
    dy = torch.randn(B, N, d)
    sd = {k: v.detach().clone().requires_grad_(True) for k, v in m.state_dict().items()}
    xo = x.clone().requires_grad_(True)
    yo = O.msa_forward(xo, sd, "", H, "vanilla")
    yo.backward(dy)
    dev = torch.device("cuda:0")
    m = m.to(dev)
    xg = x.to(dev).requires_grad_(True)
    y = m(xg)
    assert relerr(y, yo) < TOL
    y.backward(dy.to(dev))
    assert relerr(xg.grad, xo.grad) < TOL_GRAD
    for k, p in m.named_parameters():
        if sd[k].grad.abs().max() < 1e-5:      # analytically zero (a key bias shifts every score of a row alike)
            assert p.grad.abs().max().item() < 1e-4, k
            continue
        assert relerr(p.grad, sd[k].grad) < TOL_GRAD, k


@pytest.mark.parametrize("Nq,Nk,H,dh,causal,masked", [(600, 600, 2, 64, True, True), (300, 700, 2, 32, True, False),
                                                      (700, 300, 1, 128, False, True), (40, 900, 2, 64, False, False)])
def test_flash_attention_long_sequences_vs_oracle(Nq, Nk, H, dh, causal, masked):
    import kan_vit_b200 as K
    from oracle import kan_oracle as O
    K.set_precision("fp32")
    torch.manual_seed(Nq + Nk)
    B, dim = 2, 48
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
    dev = torch.device("cuda:0")
    m = m.to(dev)
    xg = x.to(dev).requires_grad_(True)
    y = m(xg, context=None if ctx is None else ctx.to(dev), mask=None if mask is None else mask.to(dev))
    assert relerr(y, yo) < TOL
    y.backward(dy.to(dev))
    assert relerr(xg.grad, xo.grad) < TOL_GRAD
    for k, p in m.named_parameters():
        assert relerr(p.grad, sd[k].grad) < TOL_GRAD, k
