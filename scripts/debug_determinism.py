"""Debug aid: run-to-run reproducibility and fp64-oracle error of the cfg-1 ViT gradients."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import kan_vit_b200 as K
from oracle import kan_oracle as O

kind = sys.argv[1] if len(sys.argv) > 1 else "efficientkan"
dev = torch.device("cuda:0")
torch.manual_seed(3)
v = K.VisionTransformer((1, 28, 28), 7, 4, 64, 2, 10, kind).to(dev)
img = torch.randn(8, 1, 28, 28, generator=torch.Generator().manual_seed(5))
dy = torch.randn(8, 10, generator=torch.Generator().manual_seed(6))

def run():
    v.zero_grad(set_to_none=True)
    y = v(img.to(dev))
    y.backward(dy.to(dev))
    return y.detach().cpu(), {k: p.grad.detach().cpu().clone() for k, p in v.named_parameters() if p.grad is not None}

sd = {k: t.detach().cpu().double().clone().requires_grad_(t.is_floating_point()) if t.is_floating_point() else t.cpu()
      for k, t in v.state_dict().items()}
yo = O.vit_forward(img.double(), sd, n_patches=7, n_blocks=4, n_heads=2, kind=kind)
yo.backward(dy.double())
y0, g0 = run()
for it in range(6):
    y, g = run()
    worst = max(((g[k].double() - sd[k].grad).abs().max().item() / max(sd[k].grad.abs().max().item(), 1e-30), k)
                for k in g if sd[k].grad.abs().max() > 1e-6)
    rep = max(((g[k] - g0[k]).abs().max().item() / max(g0[k].abs().max().item(), 1e-30), k) for k in g)
    print(f"run {it}: y err {(y.double()-yo.detach()).abs().max().item()/yo.abs().max().item():.2e} "
          f"y repro {(y-y0).abs().max().item():.2e} worst grad err {worst[0]:.2e} {worst[1]} | repro {rep[0]:.2e} {rep[1]}")
