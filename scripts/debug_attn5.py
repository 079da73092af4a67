"""timing / debugging helper: call the attention image kernels on one shape in a fresh process."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from kan_vit_b200 import _lib as L
from test_gpu_attn5 import _pack_image
B, N, H, dh = (int(a) for a in sys.argv[1:5])
lib = L.load()
dev = torch.device("cuda:0")
q, k, v = (torch.randn((B, H, N, dh)).to(torch.bfloat16).float().to(dev) for _ in range(3))
img = _pack_image(q, k, v)
out = torch.empty((B * N, H * dh), device=dev)
lse = torch.empty((B, H, N), device=dev)
st = torch.cuda.current_stream().cuda_stream
rc = lib.kv_attn_image_fwd(img.data_ptr(), None, out.data_ptr(), lse.data_ptr(), B, N, H, dh, st)
print("fwd rc", rc, lib.kv_last_cuda_error().decode() if rc else "")
torch.cuda.synchronize()
s = torch.einsum("bhid,bhjd->bhij", q, k) / dh ** 0.5
o = torch.einsum("bhij,bhjd->bhid", torch.softmax(s, -1), v).permute(0, 2, 1, 3).reshape(B * N, H * dh)
print("fwd err", ((out - o).abs().max() / o.abs().max()).item())
if len(sys.argv) > 5:
    iters = int(sys.argv[5])
    dout = torch.randn_like(out)
    T8 = (B * N + 255) // 256 * 32
    O8 = 3 * dh // 8
    dy = torch.empty((H * T8 * O8 * 128,), dtype=torch.uint8, device=dev)
    for name, fn in (("fwd", lambda: lib.kv_attn_image_fwd(img.data_ptr(), None, out.data_ptr(), lse.data_ptr(), B, N, H, dh, st)),
                     ("bwd", lambda: lib.kv_attn_image_bwd(img.data_ptr(), dout.data_ptr(), lse.data_ptr(), dy.data_ptr(), B, N, H, dh, T8, O8, st))):
        for _ in range(3):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            rc = fn()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / iters
        fl = (4.0 if name == "fwd" else 8.0) * B * H * N * N * dh
        print(f"{name}: rc {rc} {ms:.4f} ms, {fl / ms / 1e9:.1f} TFLOP/s algorithmic")
