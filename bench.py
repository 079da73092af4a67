#!/usr/bin/env python
"""bench.py -- KAN-ViT training throughput on B200 (contract in the task brief).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --steps K --warmup W    # CPU arm: the reference algorithm
                                                             # (oracle port) on the host cores

A "step" = one training iteration of the path on one synthetic batch per GPU:
zero_grad -> forward -> CrossEntropy -> backward -> (gradient all-reduce) -> Adam
(the step shape of the reference's train.py:31-40).  Workload at every N: BASELINE.json
configs[1], the ChebyKAN ViT (degree 4) on MNIST-shaped 1x28x28 input, batch 1024 per GPU
(weak scaling).  Other configs are selectable with --config for the record.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

CONFIGS = {
    # name: (chw, n_patches, n_blocks, d_hidden, n_heads, out_d, type, kan_mlp, batch/GPU, cpu sample batch)
    "cfg1": ((1, 28, 28), 7, 4, 64, 2, 10, "efficientkan", False, 64, 64),
    "cfg2": ((1, 28, 28), 7, 4, 64, 2, 10, "cheby", False, 1024, 128),
    "cfg3_fast": ((3, 32, 32), 4, 6, 192, 8, 10, "fast", False, 1024, 32),
    "cfg3_sine": ((3, 32, 32), 4, 6, 192, 8, 10, "sine", False, 1024, 32),
    "cfg4": ((3, 224, 224), 14, 12, 768, 12, 10, "efficientkan", True, 64, 1),
    "cfg4_faithful": ((3, 224, 224), 14, 12, 768, 12, 10, "efficientkan", False, 64, 2),
}
METRIC = "train_images_per_sec"
UNIT = "images/s"


def workload_name(name, batch):
    chw, npatch, nb, d, h, out_d, kind, kan_mlp, _, _ = CONFIGS[name]
    mlp = "KAN MLP" if kan_mlp else "Linear/ReLU/Linear MLP"
    return (f"{name}: KAN-ViT type={kind} {chw[0]}x{chw[1]}x{chw[2]} n_patches={npatch} d={d} depth={nb} "
            f"heads={h} {mlp}, batch {batch}/GPU, step = zero_grad+fwd+CE+bwd+Adam")


# ---------------------------------------------------------------------------------------------
# clocks: sampled with NVML from a side thread during the timed region
# ---------------------------------------------------------------------------------------------
class ClockSampler:
    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thr = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        nv = self.nv
        names = {
            nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
            nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
            nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
            nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap",
        }
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, nm in names.items():
                    if r & bit:
                        self.reasons.add(nm)
            except Exception:
                pass
            self._stop.wait(0.05)

    def start(self):
        if self.nv is not None:
            self._thr = threading.Thread(target=self._run, daemon=True)
            self._thr.start()

    def stop(self):
        self._stop.set()
        if self._thr is not None:
            self._thr.join()
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(s)}


# ---------------------------------------------------------------------------------------------
# CPU arm: the reference algorithm (oracle port) on the host cores
# ---------------------------------------------------------------------------------------------
def cpu_reference_step_fn(name, batch, seed=0):
    """Returns (step, n_images).  TEST-INFRASTRUCTURE oracle used as the reported CPU baseline."""
    from oracle import kan_oracle as O
    import kan_vit_b200 as K
    chw, npatch, nb, d, h, out_d, kind, kan_mlp, _, _ = CONFIGS[name]
    torch.manual_seed(seed)
    proto = K.VisionTransformer(chw, npatch, nb, d, h, out_d, kind, kan_mlp=kan_mlp)  # init only (CPU tensors)
    sd = {k: v.detach().clone().requires_grad_(v.is_floating_point() and k.split(".")[-1] not in ("grid", "phase"))
          for k, v in proto.state_dict().items()}
    params = [v for v in sd.values() if v.requires_grad]
    opt = torch.optim.Adam(params, lr=1e-3)
    g = torch.Generator().manual_seed(seed)
    img = torch.randn((batch,) + tuple(chw), generator=g)
    lab = torch.randint(0, out_d, (batch,), generator=g)

    def step():
        opt.zero_grad(set_to_none=True)
        y = O.vit_forward(img, sd, n_patches=npatch, n_blocks=nb, n_heads=h, kind=kind, kan_mlp=kan_mlp)
        loss = torch.nn.functional.cross_entropy(y, lab)
        loss.backward()
        opt.step()
        return float(loss)

    return step, batch


def time_cpu(name, batch, steps, warmup):
    step, n = cpu_reference_step_fn(name, batch)
    for _ in range(warmup):
        step()
    ts = []
    for _ in range(steps):
        t0 = time.perf_counter()
        step()
        ts.append(time.perf_counter() - t0)
    mean = sum(ts) / len(ts)
    return n / mean, mean * 1e3


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    name = args.config
    sample = CONFIGS[name][9]
    cores = torch.get_num_threads()
    val, ms = time_cpu(name, sample, args.steps, max(args.warmup, 1))
    line = {
        "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic", "impl": "reference",
        "config": {"workload": workload_name(name, CONFIGS[name][8]),
                   "reference_sample": f"batch {sample} per step on the host CPU"},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"{args.steps} steps of batch {sample} (same model, oracle/kan_oracle.py "
                                   "restatement of the reference, torch CPU fp32, autograd + Adam)"},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)
