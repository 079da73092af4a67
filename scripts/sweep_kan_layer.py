"""BASELINE.json config 5: the KANLinear microbenchmark sweep on one B200 (bf16 tensor-core mode).

    python scripts/sweep_kan_layer.py [--quick] [--families effkan,cheby,...] > profiles/rNN_kan_layer_sweep.jsonl

tokens {4Ki, 16Ki, 64Ki, 256Ki, 1Mi} x in=out {192, 384, 768, 1536, 3072} (+ 768<->3072) x grid {5, 10, 20} x the five
basis families (+ nn.Linear once per shape).  One JSON line per point: forward and forward+backward TFLOP/s with the
algorithmic FLOP counts of SURVEY.md section 8d (fwd 2*T*in*out*M, fwd+bwd 6*T*in*out*M), the fraction of the measured
sustained / burst bf16 peaks (MEASURED_PEAKS.json), and the NVML clock record of the timed loop.  The CPU column
comes from `bench.py --impl reference --config cfg5 --sweep-cpu`."""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "scripts"))
import torch

import kan_vit_b200 as K
from kan_vit_b200 import ops
from kan_vit_b200._dispatch import make_spec, slots_of
from bench import ClockSampler, load_peaks
from bench_kan_layer import make


def _fast_reset(self):
    # the least-squares fit of KANLinear.reset_parameters takes seconds per layer at 3072^2 x G=20 and does not change
    # a single launch: same shapes, weights of the same scale
    with torch.no_grad():
        torch.nn.init.kaiming_uniform_(self.base_weight, a=5 ** 0.5 * self.scale_base)
        self.spline_weight.normal_(0.0, 0.02)
        if self.enable_standalone_scale_spline:
            torch.nn.init.kaiming_uniform_(self.spline_scaler, a=5 ** 0.5 * self.scale_spline)


def run(family, T, i, o, grid, peaks):
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    m = make(family, i, o, grid).to(dev)
    spec = make_spec([m])
    x = torch.randn(T, i, device=dev, requires_grad=True)
    dy = torch.randn(T, o, device=dev)
    flops = 2.0 * T * i * o * spec.M

    def step():
        x.grad = None
        ops.kan_apply(x, spec, [slots_of(m)]).backward(dy)

    t0 = time.perf_counter()
    step()
    torch.cuda.synchronize()
    first = time.perf_counter() - t0
    iters = max(1, min(5, int(0.5 / max(first, 1e-3))))
    if first < 2.0:
        step()
        torch.cuda.synchronize()
    clocks = ClockSampler(0)
    clocks.start()
    ops.PROFILE = {}
    for _ in range(iters):
        step()
    torch.cuda.synchronize()
    clk = clocks.stop()
    prof, ops.PROFILE = ops.PROFILE, None
    med = {}
    for name, evs in prof.items():
        ms = sorted(s.elapsed_time(e) for s, e, *_ in evs)
        med[name] = ms[len(ms) // 2]
    gemm = [med.get(k, 0.0) for k in ("kv_kan_fwd", "kv_kan_bwd_input", "kv_kan_bwd_weight")]
    allk = sum(med.values())
    fb = 3 * flops / (sum(gemm) * 1e-3) / 1e12
    return {"family": family, "T": T, "in": i, "out": o, "grid": grid, "M": spec.M, "iters": iters,
            "fwd_ms": round(gemm[0], 4), "bwd_input_ms": round(gemm[1], 4), "bwd_weight_ms": round(gemm[2], 4),
            "all_kernels_ms": round(allk, 4),
            "fwd_tflops": round(flops / (gemm[0] * 1e-3) / 1e12, 1),
            "bwd_input_tflops": round(flops / (gemm[1] * 1e-3) / 1e12, 1),
            "bwd_weight_tflops": round(flops / (gemm[2] * 1e-3) / 1e12, 1),
            "fwdbwd_tflops_gemm_kernels": round(fb, 1),
            "fwdbwd_tflops_incl_pack_and_layout": round(3 * flops / (allk * 1e-3) / 1e12, 1),
            "frac_sustained": round(fb / peaks["bf16_tflops_sustained"], 3), "frac_burst": round(fb / peaks["bf16_tflops"], 3),
            "clocks": clk}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--families", default="effkan,cheby,fastkan,sine,fourier,linear")
    ap.add_argument("--tokens", default="4096,16384,65536,262144,1048576")
    ap.add_argument("--grids", default="5,10,20")
    a = ap.parse_args()
    K.set_precision("bf16")
    K.KANLinear.reset_parameters = _fast_reset
    peaks = load_peaks()
    fams = a.families.split(",")
    shapes = [(192, 192), (384, 384), (768, 768), (1536, 1536), (3072, 3072), (768, 3072), (3072, 768)]
    if a.quick:
        cases = [(f, 16384, 768, 768, 5) for f in fams]
    else:
        cases = []
        for f in fams:
            for grid in ([5] if f == "linear" else [int(g) for g in a.grids.split(",")]):
                for T in (int(t) for t in a.tokens.split(",")):
                    cases += [(f, T, i, o, grid) for i, o in shapes]
    print(json.dumps({"peaks": peaks, "points": len(cases)}), flush=True)
    for c in cases:
        try:
            print(json.dumps(run(*c, peaks)), flush=True)
        except Exception as e:  # a shape a kernel refuses is part of the record
            print(json.dumps({"family": c[0], "T": c[1], "in": c[2], "out": c[3], "grid": c[4], "error": str(e)[:200]}),
                  flush=True)
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
