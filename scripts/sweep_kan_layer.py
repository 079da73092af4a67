"""BASELINE.json config 5: KAN-layer microbenchmark sweep on one B200 (bf16 tensor-core mode).

    python scripts/sweep_kan_layer.py [--quick] > profiles/rNN_kan_layer_sweep.jsonl

tokens x in/out x grid across the five basis families (+ nn.Linear); forward and forward+backward TFLOP/s with
the algorithmic FLOP counts of SURVEY.md section 8d (fwd 2*T*in*out*M, fwd+bwd 6*T*in*out*M).  Gradient kernels of
the families without a tensor-core gradient kernel run on the exact fp32 kernels."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import kan_vit_b200 as K
from kan_vit_b200 import ops
from kan_vit_b200._dispatch import make_spec, slots_of
from bench_kan_layer import make


def run(family, T, i, o, grid, iters=5):
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    m = make(family, i, o, grid).to(dev)
    spec = make_spec([m])
    x = torch.randn(T, i, device=dev, requires_grad=True)
    dy = torch.randn(T, o, device=dev)

    def step():
        ops.kan_apply(x, spec, [slots_of(m)]).backward(dy)

    for _ in range(2):
        step()
    torch.cuda.synchronize()
    ops.PROFILE = {}
    for _ in range(iters):
        step()
    torch.cuda.synchronize()
    prof, ops.PROFILE = ops.PROFILE, None
    med = {}
    for name, evs in prof.items():
        ms = sorted(s.elapsed_time(e) for s, e, *_ in evs)
        med[name] = ms[len(ms) // 2]
    flops = 2.0 * T * i * o * spec.M
    gemm = [med.get(k, 0.0) for k in ("kv_kan_fwd", "kv_kan_bwd_input", "kv_kan_bwd_weight")]
    allk = sum(med.values())
    return {"family": family, "T": T, "in": i, "out": o, "grid": grid, "M": spec.M,
            "fwd_ms": round(gemm[0], 4), "bwd_input_ms": round(gemm[1], 4), "bwd_weight_ms": round(gemm[2], 4),
            "all_kernels_ms": round(allk, 4),
            "fwd_tflops": round(flops / (gemm[0] * 1e-3) / 1e12, 1),
            "fwdbwd_tflops_gemm_kernels": round(3 * flops / (sum(gemm) * 1e-3) / 1e12, 1),
            "fwdbwd_tflops_incl_pack_and_layout": round(3 * flops / (allk * 1e-3) / 1e12, 1)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true")
    a = ap.parse_args()
    K.set_precision("bf16")
    fams = ["effkan", "cheby", "fastkan", "sine", "fourier", "linear"]
    if a.quick:
        cases = [(f, 16384, 768, 768, 5) for f in fams]
    else:
        cases = []
        for f in fams:
            cases += [(f, 65536, 768, 768, 5), (f, 16384, 192, 192, 5), (f, 4096, 3072, 3072, 5)]
        cases += [("effkan", T, 768, 3072, 5) for T in (4096, 16384, 65536, 262144)]
        cases += [("effkan", 65536, 3072, 768, 5), ("effkan", 65536, 1536, 1536, 5), ("effkan", 65536, 384, 384, 5),
                  ("effkan", 1048576, 192, 192, 5), ("effkan", 65536, 768, 768, 10), ("effkan", 65536, 768, 768, 20),
                  ("cheby", 65536, 768, 768, 10), ("fastkan", 65536, 768, 768, 20), ("sine", 65536, 768, 768, 20),
                  ("fourier", 65536, 768, 768, 10)]
    for c in cases:
        try:
            print(json.dumps(run(*c)), flush=True)
        except Exception as e:  # a shape a kernel refuses is part of the record
            print(json.dumps({"family": c[0], "T": c[1], "in": c[2], "out": c[3], "grid": c[4], "error": str(e)[:200]}),
                  flush=True)
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
