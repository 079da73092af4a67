"""Timing experiment: SM clock and board power while ONE KAN-layer kernel runs back to back for a few seconds.
    python scripts/power_probe.py [--inf 768] [--outf 3072] [--T 66192] [--what fwd|fwdbwd] [--zeros] [--seconds 4]
Prints one JSON line: TFLOP/s of the loop, median / min SM clock, median power, throttle reasons seen."""
import argparse, json, os, sys, threading, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import pynvml
import kan_vit_b200 as K
from kan_vit_b200 import ops
from kan_vit_b200._dispatch import make_spec, slots_of

ap = argparse.ArgumentParser()
ap.add_argument("--inf", type=int, default=768)
ap.add_argument("--outf", type=int, default=3072)
ap.add_argument("--T", type=int, default=66192)
ap.add_argument("--what", default="fwd")
ap.add_argument("--family", default="effkan")
ap.add_argument("--zeros", action="store_true")
ap.add_argument("--seconds", type=float, default=4.0)
a = ap.parse_args()
dev = torch.device("cuda:0")
torch.manual_seed(0)
m = (torch.nn.Linear(a.inf, a.outf) if a.family == "linear" else K.KANLinear(a.inf, a.outf)).to(dev)
spec = make_spec([m])
K.set_precision("bf16")
x = (torch.zeros if a.zeros else torch.randn)(a.T, a.inf, device=dev).requires_grad_(a.what != "fwd")
dy = torch.randn(a.T, a.outf, device=dev)

def step():
    if a.what == "fwd":
        with torch.no_grad():
            ops.kan_apply(x, spec, [slots_of(m)])
    else:
        ops.kan_apply(x, spec, [slots_of(m)]).backward(dy)

for _ in range(3):
    step()
torch.cuda.synchronize()
pynvml.nvmlInit()
h = pynvml.nvmlDeviceGetHandleByIndex(0)
samples, stop = [], threading.Event()
def sampler():
    while not stop.is_set():
        samples.append((pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM), pynvml.nvmlDeviceGetPowerUsage(h) / 1000.0,
                        pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h)))
        time.sleep(0.05)
th = threading.Thread(target=sampler); th.start()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
n = 0
t0 = time.time(); e0.record()
while time.time() - t0 < a.seconds:
    for _ in range(20):
        step()
    n += 20
    torch.cuda.synchronize()
e1.record(); torch.cuda.synchronize()
stop.set(); th.join()
ms = e0.elapsed_time(e1) / n
flops = (2 if a.what == "fwd" else 6) * a.T * a.inf * a.outf * spec.M
clk = sorted(s[0] for s in samples[len(samples) // 4:]); pw = sorted(s[1] for s in samples[len(samples) // 4:])
reasons = 0
for s in samples: reasons |= s[2]
print(json.dumps({"family": a.family, "in": a.inf, "out": a.outf, "what": a.what, "zeros": a.zeros, "ms": round(ms, 4),
                  "tflops": round(flops / ms / 1e9, 1), "sm_mhz_median": clk[len(clk) // 2], "sm_mhz_min": clk[0],
                  "power_w_median": round(pw[len(pw) // 2], 1), "power_limit_w": pynvml.nvmlDeviceGetPowerManagementLimit(h) / 1000.0,
                  "throttle_reasons_mask": hex(reasons), "samples": len(samples)}))
