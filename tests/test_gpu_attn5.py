"""tcgen05 attention kernels (csrc/attn_tc5.cu) called directly through the C ABI on a hand-packed attention image.

Reference: softmax(q k^T / sqrt(d_h)) v (attention.py:199-200) and its autograd gradient, evaluated by torch in fp32 on
the bf16-rounded q, k, v (the image holds bf16).  Tolerance 2e-2 norm-wise (BASELINE.json north_star, bf16 paths).
Shapes: BASELINE config 4 (197 tokens, d_h 64), config 1/2 (50, 32), config 3 (17, 24), the tile edges (128 / 129 / 256
tokens) and the smallest head."""
import ctypes as C

import pytest
import torch

from conftest import relerr

pytestmark = pytest.mark.gpu
TOL = 2e-2


def _pack_image(q, k, v):
    """q, k, v [B, H, N, dh] fp32 -> uint8 image [b][h][part][n8][d8][8][8] bf16, zero padded."""
    B, H, N, dh = q.shape
    NP, DP = (N + 15) // 16 * 16, (dh + 15) // 16 * 16
    x = torch.zeros((B, H, 3, NP, DP), dtype=torch.float32, device=q.device)
    x[:, :, 0, :N, :dh], x[:, :, 1, :N, :dh], x[:, :, 2, :N, :dh] = q, k, v
    img = x.view(B, H, 3, NP // 8, 8, DP // 8, 8).permute(0, 1, 2, 3, 5, 4, 6).contiguous().to(torch.bfloat16)
    return img.view(torch.uint8).reshape(-1)


def _unpack_dy(dyimg, B, N, H, dh, T8, O8):
    """uint8 dY image [H][T8][O8][8][8] bf16 -> (dq, dk, dv) [B, H, N, dh] fp32, plus the padding for a zero check."""
    t = dyimg.view(torch.bfloat16).view(H, T8, O8, 8, 8).permute(0, 1, 3, 2, 4).reshape(H, T8 * 8, O8 * 8).float()
    body = t[:, :B * N, :3 * dh].reshape(H, B, N, 3, dh).permute(3, 1, 0, 2, 4)
    pad = torch.cat([t[:, B * N:, :].reshape(-1), t[:, :, 3 * dh:].reshape(-1)])
    return body[0], body[1], body[2], pad


SHAPES = [(3, 197, 2, 64), (2, 50, 2, 32), (2, 17, 8, 24), (1, 5, 2, 8), (2, 256, 1, 64), (2, 129, 3, 16),
          (1, 128, 2, 48), (150, 197, 1, 64)]


@pytest.mark.parametrize("B,N,H,dh", SHAPES)
def test_attention_image_forward_backward(B, N, H, dh):
    from kan_vit_b200 import _lib as L
    lib = L.load()
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(B * 1000 + N)
    q, k, v = ((torch.randn((B, H, N, dh), generator=g) * sc).to(torch.bfloat16).float().to(dev) for sc in (1.5, 1.0, 1.0))
    res = torch.randn((B * N, H * dh), generator=g).to(dev)
    dout = torch.randn((B * N, H * dh), generator=g).to(dev)
    img = _pack_image(q, k, v)
    assert img.numel() == lib.kv_attn_image_bytes(B, N, H, dh)
    out = torch.empty((B * N, H * dh), device=dev)
    lse = torch.empty((B, H, N), device=dev)
    st = torch.cuda.current_stream().cuda_stream
    L.check(lib.kv_attn_image_fwd(img.data_ptr(), res.data_ptr(), out.data_ptr(), lse.data_ptr(), B, N, H, dh, st), "fwd")
    torch.cuda.synchronize()
    # reference
    qr, kr, vr = (t.clone().requires_grad_(True) for t in (q, k, v))
    s = torch.einsum("bhid,bhjd->bhij", qr, kr) / dh ** 0.5
    o = torch.einsum("bhij,bhjd->bhid", torch.softmax(s, -1), vr)           # [B,H,N,dh]
    o2 = o.permute(0, 2, 1, 3).reshape(B * N, H * dh)
    assert relerr(out - res, o2) < TOL
    assert relerr(lse, torch.logsumexp(s, -1)) < 1e-3
    o2.backward(dout)
    # backward: dY image with padding on both axes
    T8 = (B * N + 255) // 256 * 32
    O8 = max(3 * dh // 8, 8) + (1 if dh == 24 else 0)
    dy = torch.full((H * T8 * O8 * 128,), 0x7F, dtype=torch.uint8, device=dev)   # poison: padding must come back zero
    L.check(lib.kv_attn_image_bwd(img.data_ptr(), dout.data_ptr(), lse.data_ptr(), dy.data_ptr(), B, N, H, dh, T8, O8, st),
            "bwd")
    torch.cuda.synchronize()
    dq, dk, dv, pad = _unpack_dy(dy, B, N, H, dh, T8, O8)
    assert pad.numel() == 0 or pad.abs().max().item() == 0.0
    assert relerr(dv, vr.grad) < TOL
    assert relerr(dq, qr.grad) < TOL
    assert relerr(dk, kr.grad) < TOL


def test_attention_image_shape_probe():
    from kan_vit_b200 import _lib as L
    lib = L.load()
    assert lib.kv_attn_image_fwd(None, None, None, None, 0, 197, 12, 64, None) == 0
    assert lib.kv_attn_image_fwd(None, None, None, None, 0, 257, 12, 64, None) != 0     # more than two query tiles
    assert lib.kv_attn_image_fwd(None, None, None, None, 0, 50, 2, 20, None) != 0       # d_head % 8 != 0
    assert lib.kv_attn_image_fwd(None, None, None, None, 0, 50, 2, 128, None) != 0
