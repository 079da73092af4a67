"""MSA microbenchmark: the per-head KAN q/k/v maps + attention of one block (attention.py:181-202), forward + backward,
bf16 tensor-core mode, at the default bench's shape (336 images x 197 tokens, d = 768, 12 heads).
    python scripts/bench_msa.py [--B 336] [--N 197] [--d 768] [--heads 12] [--type efficientkan] [--iters 5]
Prints one JSON line with the median device time per entry point (CUDA events around every C-ABI call)."""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import kan_vit_b200 as K
from kan_vit_b200 import ops

ap = argparse.ArgumentParser()
ap.add_argument("--B", type=int, default=336)
ap.add_argument("--N", type=int, default=197)
ap.add_argument("--d", type=int, default=768)
ap.add_argument("--heads", type=int, default=12)
ap.add_argument("--type", default="efficientkan")
ap.add_argument("--iters", type=int, default=5)
a = ap.parse_args()
dev = torch.device("cuda:0")
torch.manual_seed(0)
K.set_precision("bf16")
m = K.MSA(a.d, n_heads=a.heads, type=a.type).to(dev)
x = torch.randn(a.B, a.N, a.d, device=dev, requires_grad=True)
dy = torch.randn(a.B, a.N, a.d, device=dev)


def step():
    x.grad = None
    m(x).backward(dy)


for _ in range(3):
    step()
torch.cuda.synchronize()
ops.PROFILE = {}
for _ in range(a.iters):
    step()
torch.cuda.synchronize()
prof, ops.PROFILE = ops.PROFILE, None
out = {"B": a.B, "N": a.N, "d": a.d, "heads": a.heads, "type": a.type}
for name, evs in prof.items():
    ms = sorted(s.elapsed_time(e) for s, e, *_ in evs)
    out[name] = round(ms[len(ms) // 2], 4)
out["total_ms"] = round(sum(v for k, v in out.items() if k.startswith("kv_")), 4)
print(json.dumps(out))
