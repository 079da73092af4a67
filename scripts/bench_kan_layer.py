"""KAN-layer microbenchmark (BASELINE.json config 5): fused layer TFLOP/s on one B200.

    python scripts/bench_kan_layer.py [--family effkan] [--T 25216] [--inf 768] [--outf 3072] [--grid 5]
                                      [--precision bf16] [--what fwd|fwdbwd]
Algorithmic FLOPs: fwd 2*T*in*out*M, fwd+bwd 6*T*in*out*M."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import kan_vit_b200 as K
from kan_vit_b200 import ops
from kan_vit_b200._dispatch import make_spec, slots_of


def make(family, i, o, grid):
    return {
        "effkan": lambda: K.KANLinear(i, o, grid_size=grid),
        "cheby": lambda: K.ChebyKANLayer(i, o, grid),
        "fastkan": lambda: K.FastKANLayer(i, o, num_grids=grid),
        "sine": lambda: K.SineKANLayer(i, o, grid_size=grid),
        "fourier": lambda: K.NaiveFourierKANLayer(i, o, grid),
        "linear": lambda: torch.nn.Linear(i, o),
    }[family]()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--family", default="effkan")
    ap.add_argument("--T", type=int, default=25216)
    ap.add_argument("--inf", type=int, default=768)
    ap.add_argument("--outf", type=int, default=3072)
    ap.add_argument("--grid", type=int, default=5)
    ap.add_argument("--precision", default="bf16")
    ap.add_argument("--what", default="fwd")
    ap.add_argument("--iters", type=int, default=10)
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    m = make(a.family, a.inf, a.outf, a.grid).to(dev)
    spec = make_spec([m])
    x = torch.randn(a.T, a.inf, device=dev, requires_grad=(a.what != "fwd"))
    dy = torch.randn(a.T, a.outf, device=dev)
    K.set_precision(a.precision)

    def step():
        if a.what == "fwd":
            with torch.no_grad():
                return ops.kan_apply(x, spec, [slots_of(m)])
        y = ops.kan_apply(x, spec, [slots_of(m)])
        y.backward(dy)
        return y

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    ops.PROFILE = {}
    for _ in range(a.iters):
        step()
    torch.cuda.synchronize()
    prof, ops.PROFILE = ops.PROFILE, None
    out = {"family": a.family, "T": a.T, "in": a.inf, "out": a.outf, "M": spec.M, "precision": a.precision}
    for name, evs in prof.items():
        ms = sorted(s.elapsed_time(e) for s, e, *_ in evs)
        med = ms[len(ms) // 2]
        w = evs[0][2]
        out[name] = {"ms": round(med, 4), "tflops": round(w / (med * 1e-3) / 1e12, 2) if w and name.startswith("kv_kan_") and name != "kv_kan_dy_image" and name != "kv_kan_pack" else None}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
