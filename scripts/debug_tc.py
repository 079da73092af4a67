import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import kan_vit_b200 as K
from kan_vit_b200 import ops
from kan_vit_b200._dispatch import make_spec, slots_of
dev = torch.device("cuda:0")
def run(fam, T, i, o):
    torch.manual_seed(0)
    m = {"effkan": lambda: K.KANLinear(i, o), "cheby": lambda: K.ChebyKANLayer(i, o, 4)}[fam]().to(dev)
    x = torch.randn(T, i, device=dev)
    spec = make_spec([m])
    K.set_precision("bf16"); y16 = ops.kan_apply(x, spec, [slots_of(m)])
    K.set_precision("fp32"); y32 = ops.kan_apply(x, spec, [slots_of(m)])
    e = (y16 - y32).abs()
    rowerr = e.max(dim=1).values; colerr = e.max(dim=0).values
    print(fam, T, i, o, "rel", (e.max() / y32.abs().max()).item(),
          "bad rows", (rowerr > 0.05 * y32.abs().max()).sum().item(), "bad cols", (colerr > 0.05 * y32.abs().max()).sum().item())
for cfg in [("effkan", 128, 16, 16), ("effkan", 128, 32, 16), ("effkan", 128, 48, 16), ("effkan", 128, 64, 16), ("effkan", 128, 80, 16),
            ("effkan", 128, 48, 528), ("effkan", 128, 64, 528), ("cheby", 128, 128, 16), ("cheby", 128, 96, 16), ("effkan", 300, 80, 528)]:
    run(*cfg)
