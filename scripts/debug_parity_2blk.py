"""A/B helper: bf16 errors of the 2-block config-4 ViT against the fp32 oracle (KV_NO_FUSED_ATTN=1 = old attention route)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import test_big_parity as T
name = sys.argv[1] if len(sys.argv) > 1 else "big_vit_cfg4_2blk_kanmlp_B2"
m, fwd, x, seed = T._case(name)
sd = T._golden_state(m, seed)
m.load_state_dict(sd, strict=False)
m = m.to("cuda:0")
with torch.no_grad():
    shape = fwd(x, sd).shape
dy = T._dy_for(name, shape, seed)
y32, gx32, g32 = T._oracle_run(fwd, x, sd, dy, torch.float32)
y, gx, grads = T._gpu_run(m, x, dy, "bf16")
e = T._errors(y, gx, grads, y32, gx32, g32)
top = sorted(e.items(), key=lambda kv: -kv[1])[:12]
print("fused" if os.environ.get("KV_NO_FUSED_ATTN") != "1" else "old route", "y", e["y"], "gx", e["gx"])
for k, v in top:
    print(f"  {v:.4f} {k}")
