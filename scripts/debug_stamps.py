import os, sys
os.environ["KV_DEBUG_BITS"] = sys.argv[1] if len(sys.argv) > 1 else "32"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import kan_vit_b200 as K
from kan_vit_b200 import ops
from kan_vit_b200._dispatch import make_spec, slots_of
dev = torch.device("cuda:0")
INF = int(sys.argv[2]) if len(sys.argv) > 2 else 768
OUTF = int(sys.argv[3]) if len(sys.argv) > 3 else 3072
m = K.KANLinear(INF, OUTF).to(dev)
x = torch.randn(int(os.environ.get("KV_T", "25216")), INF, device=dev)
spec = make_spec([m]); K.set_precision("bf16")
for _ in range(2):
    y = ops.kan_apply(x, spec, [slots_of(m)])
torch.cuda.synchronize()
for rank in range(2):
    d = y[rank * 128: rank * 128 + 2, :512].contiguous().view(torch.int32).cpu().view(4, 64, 4).long() & 0xffffffff
    t0 = int(d[1, 0, 0])
    f = lambda v: (int(v) - t0) & 0xffffffff
    print("CTA rank", rank, "roles: 0=mma(stage: before fullB wait, after, before fullA wait (part start), -) 1=loader(u: emptyB ok, issued) 2=producer(chunk: emptyA ok, arrived) 3=relay(u: fullB, arrived)")
    print("tile: after set-up sync", f(d[2, 61, 1]), "producers start", f(d[2, 62, 0]), "producers done", f(d[2, 62, 1]),
          "accumulator complete", f(d[2, 62, 2]), "epilogue done (warp 2)", f(d[2, 62, 3]), "all warps done", f(d[2, 61, 2]))
    for blk in (60, 59):
        print("  epilogue column block", 60 - blk, "of warp 2: before tcgen05.ld", f(d[2, blk, 0]), "loaded", f(d[2, blk, 1]),
              "staged in smem", f(d[2, blk, 2]), "stores issued", f(d[2, blk, 3]))
    if os.environ.get("KV_STAMPS_SHORT"):
        continue
    for u in list(range(0, 64)):
        print(u, "mma", [f(d[0, u, k]) for k in range(4)], "load", [f(d[1, u, k]) for k in range(2)],
              "prod", [f(d[2, u, k]) for k in range(2)], "relay", [f(d[3, u, k]) for k in range(2)])
