#!/usr/bin/env python
"""bench.py -- KAN-ViT training throughput on B200 (contract in the task brief).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --steps K --warmup W    # CPU arm: the reference algorithm
                                                             # (oracle port) on the host cores

A "step" = one training iteration of the path on one synthetic batch per GPU:
zero_grad -> forward -> CrossEntropy -> backward -> (gradient all-reduce) -> Adam
(the step shape of the reference's train.py:31-40).  Workload at every N: the configuration
BASELINE.json's metric ("train images/sec at 1/2/4/8 B200") is quoted on, configs[3]: the
ViT-B/16-shaped KAN-ViT (efficient-kan MLP, 224x224, 197 tokens, dim 768, depth 12), bf16
tensor-core arithmetic, data-parallel, 336 images per GPU (weak scaling; 336 x 197 token rows = 259
row blocks of 256, which the three layer shapes turn into exactly 7 / 21 / 42 waves of CTA pairs on the
74 pair slots of a B200 -- a multiple of 48 images does that).  It fits one GPU, so it
is also the N=1 workload.  The other configs (cfg1 .. cfg3, parity-test shapes) are selectable
with --config for the record; cfg1/cfg2 default to the exact fp32 kernels.
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
    "cfg2": ((1, 28, 28), 7, 4, 64, 2, 10, "cheby", False, 1024, 1024),
    "cfg3_fast": ((3, 32, 32), 4, 6, 192, 8, 10, "fast", False, 1024, 32),
    "cfg3_sine": ((3, 32, 32), 4, 6, 192, 8, 10, "sine", False, 1024, 32),
    "cfg4": ((3, 224, 224), 14, 12, 768, 12, 10, "efficientkan", True, 336, 4),
    "cfg4_faithful": ((3, 224, 224), 14, 12, 768, 12, 10, "efficientkan", False, 336, 2),
}
DEFAULT_PRECISION = {"cfg1": "fp32", "cfg2": "fp32", "cfg3_fast": "bf16", "cfg3_sine": "bf16", "cfg4": "bf16",
                     "cfg4_faithful": "bf16"}
CPU_TIME_BUDGET_S = 120.0  # the reference arm stops early (and says so) once its timed steps have used this
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
def cpu_reference_step_fn(name, batch, seed=0, optimizer=True):
    """Returns (step, n_images).  TEST-INFRASTRUCTURE oracle used as the reported CPU baseline.  Nothing of the
    product package is imported here: the initial state comes from oracle.kan_oracle.init_state."""
    from oracle import kan_oracle as O
    chw, npatch, nb, d, h, out_d, kind, kan_mlp, _, _ = CONFIGS[name]
    sd = O.init_state(chw, npatch, nb, d, h, out_d, kind, kan_mlp=kan_mlp, seed=seed)
    sd = {k: v.requires_grad_(v.is_floating_point() and k.split(".")[-1] not in ("grid", "phase", "arange"))
          for k, v in sd.items()}
    params = [v for v in sd.values() if v.requires_grad]
    opt = torch.optim.Adam(params, lr=1e-3) if optimizer else None
    g = torch.Generator().manual_seed(seed)
    img = torch.randn((batch,) + tuple(chw), generator=g)
    lab = torch.randint(0, out_d, (batch,), generator=g)

    def step():
        for p in params:
            p.grad = None
        y = O.vit_forward(img, sd, n_patches=npatch, n_blocks=nb, n_heads=h, kind=kind, kan_mlp=kan_mlp)
        loss = torch.nn.functional.cross_entropy(y, lab)
        loss.backward()
        if opt is not None:
            opt.step()
        return loss.item()

    return step, batch


def time_cpu(name, batch, steps, warmup, budget_s=None, optimizer=True):
    """(images/s, ms/step, steps actually timed).  With budget_s the loop stops once the timed steps
    have used that much wall time (the ViT-B-shaped config takes seconds per image on the host)."""
    step, n = cpu_reference_step_fn(name, batch, optimizer=optimizer)
    for _ in range(warmup):
        step()
    ts = []
    for _ in range(steps):
        t0 = time.perf_counter()
        step()
        ts.append(time.perf_counter() - t0)
        if budget_s is not None and sum(ts) > budget_s:
            break
    mean = sum(ts) / len(ts)
    return n / mean, mean * 1e3, len(ts)


def host_threads():
    """All the host threads the CPU arm can use: torchrun exports OMP_NUM_THREADS=1 to its workers, which would
    pin the arm to one core at N > 1, so the count is set explicitly."""
    try:
        n = len(os.sched_getaffinity(0))
    except AttributeError:
        n = os.cpu_count() or 1
    torch.set_num_threads(n)
    return torch.get_num_threads()


def time_config1(cores, steps=3):
    """BASELINE.json configs[0] exactly as stated: ViKANformer efficient-kan (grid 5, order 3), 1x28x28, patch 7,
    dim 64, depth 4, batch 64, forward + CrossEntropy + backward on the CPU (no optimizer), with all host threads
    and with one (the reference path is dispatch-bound: SURVEY.md section 6)."""
    out = {"workload": workload_name("cfg1", 64).replace("+Adam", "") + " (fwd+CE+bwd, BASELINE configs[0])"}
    v, ms, n = time_cpu("cfg1", 64, steps, 1, 40.0, optimizer=False)
    out.update({"value": v, "unit": UNIT, "cores": cores, "ms_per_step": ms, "steps": n})
    torch.set_num_threads(1)
    v1, ms1, n1 = time_cpu("cfg1", 64, steps, 1, 40.0, optimizer=False)
    torch.set_num_threads(cores)
    out["one_thread"] = {"value": v1, "unit": UNIT, "cores": 1, "ms_per_step": ms1, "steps": n1}
    return out


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    name = args.config
    sample = CONFIGS[name][9]
    cores = host_threads()
    print(f"[bench reference arm] {cores} host threads (torch.get_num_threads), sample batch {sample}",
          file=sys.stderr, flush=True)
    warm = 1
    val, ms, done = time_cpu(name, sample, args.steps, warm, CPU_TIME_BUDGET_S)
    cfg1 = time_config1(cores)
    line = {
        "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": done,
        "warmup": warm, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic", "impl": "reference",
        "config": {"workload": workload_name(name, CONFIGS[name][8]),
                   "reference_sample": f"batch {sample} per step on the host CPU, {done} timed steps "
                                       f"(stops after {CPU_TIME_BUDGET_S:.0f} s of timed steps)"},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"{done} steps of batch {sample} (same model and step, oracle/kan_oracle.py "
                                   "restatement of the reference, torch CPU fp32, autograd + Adam)",
                         "config1": cfg1},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), file=_OUT, flush=True)


# ---------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------
def load_peaks():
    """Roofline denominators: MEASURED_PEAKS.json (driver-written) or the profiling guide's fallback."""
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return {"hbm_gbs": d["hbm_gbs"], "bf16_tflops": d["bf16_tflops"],
                "bf16_tflops_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]), "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


# kernels whose work is counted in FLOPs (tensor roofline); the rest are counted in bytes (HBM roofline)
FLOP_KERNELS = {"kv_kan_fwd", "kv_kan_bwd_input", "kv_kan_bwd_weight", "kv_attn_fwd", "kv_attn_bwd",
                "kv_attn_fwd_bf16", "kv_attn_bwd_bf16"}


def profile_kernels(step_fn, n_steps):
    """Per-entry-point device time and algorithmic work of `n_steps` steps, measured with CUDA
    events recorded on the launching stream around every C-ABI call (kan_vit_b200.ops._call)."""
    from kan_vit_b200 import ops
    ops.PROFILE = {}
    for _ in range(n_steps):
        step_fn()
    torch.cuda.synchronize()
    prof, ops.PROFILE = ops.PROFILE, None
    out = {}
    by_shape = {}
    for name, evs in prof.items():
        ms = sum(s.elapsed_time(e) for s, e, _, _ in evs)
        work = sum(w for _, _, w, _ in evs)
        out[name] = {"calls": len(evs) // n_steps, "ms_per_step": ms / n_steps, "work_per_step": work / n_steps,
                     "ms_per_call": ms / len(evs), "work_per_call": work / len(evs)}
        for s, e, w, tag in evs:
            if tag is not None:
                a = by_shape.setdefault(f"{name} {tag}", [0.0, 0.0, 0])
                a[0] += s.elapsed_time(e)
                a[1] += w
                a[2] += 1
    out["_by_shape"] = {k: {"calls": v[2] // n_steps, "ms_per_call": v[0] / v[2],
                            "tflops": v[1] / v[2] / (v[0] / v[2] * 1e-3) / 1e12}
                        for k, v in by_shape.items()}
    return out


def _progress(msg):
    if os.environ.get("KV_VERBOSE") == "1":
        print(f"[bench +{time.perf_counter() - _T0:7.1f}s] {msg}", file=sys.stderr, flush=True)


_T0 = time.perf_counter()


def run_gpu_arm(args):
    import torch.distributed as dist
    import kan_vit_b200 as K
    from kan_vit_b200 import _lib as L
    from kan_vit_b200.trainer import TrainStep

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the GPU arm has no CPU fallback; use --impl reference)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lib = L.load()

    name = args.config
    chw, npatch, nb, d, h, out_d, kind, kan_mlp, batch, cpu_batch = CONFIGS[name]
    if args.batch:
        batch = args.batch
    K.set_precision(args.precision)
    if os.environ.get("KV_FAST_INIT") == "1":
        # profiler runs only: skip the least-squares fit of KANLinear.reset_parameters (minutes of host
        # LAPACK work under ncu); same shapes, same launches, weights of the same scale
        def _fast_reset(self):
            with torch.no_grad():
                torch.nn.init.kaiming_uniform_(self.base_weight, a=5 ** 0.5 * self.scale_base)
                self.spline_weight.normal_(0.0, 0.02)
                if self.enable_standalone_scale_spline:
                    torch.nn.init.kaiming_uniform_(self.spline_scaler, a=5 ** 0.5 * self.scale_spline)
        K.KANLinear.reset_parameters = _fast_reset
    torch.manual_seed(0)  # identical replicas on every rank
    model = K.VisionTransformer(chw, npatch, nb, d, h, out_d, kind, kan_mlp=kan_mlp).to(dev)
    _progress("model on device")
    if world > 1:
        for p in model.parameters():
            dist.broadcast(p.data, 0)
    step = TrainStep(model, lr=1e-3, cuda_graph=args.cuda_graph)
    g = torch.Generator().manual_seed(1000 + rank)  # every rank its own shard of synthetic data
    n_host = 4  # rotating pinned host batches for the end-to-end leg
    host_img = [torch.randn((batch,) + tuple(chw), generator=g).pin_memory() for _ in range(n_host)]
    host_lab = [torch.randint(0, out_d, (batch,), generator=g).pin_memory() for _ in range(n_host)]
    dev_img = [t.to(dev) for t in host_img]
    dev_lab = [t.to(dev) for t in host_lab]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        """ms for `steps` calls of fn(i): barrier+sync, CUDA events on the launching stream, MAX over ranks."""
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            fn(i)
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    # KV_NCU_RANGE=1: bracket the first timed step with cudaProfilerStart/Stop so that
    # `ncu --profile-from-start off` captures exactly one step of this command (profiles/README.md)
    ncu_range = os.environ.get("KV_NCU_RANGE") == "1"
    state = {"timed": False, "done": False}

    def resident_step(i):
        mark = ncu_range and state["timed"] and not state["done"]
        if mark:
            torch.cuda.synchronize()
            torch.cuda.profiler.start()
        step(dev_img[i % n_host], dev_lab[i % n_host])
        if mark:
            torch.cuda.synchronize()
            torch.cuda.profiler.stop()
            state["done"] = True

    loss_host = [torch.zeros((), dtype=torch.float32).pin_memory() for _ in range(2)]
    loss_evt = [torch.cuda.Event() for _ in range(2)]
    e2e_losses = []

    from kan_vit_b200.trainer import BatchPrefetcher
    feeder = BatchPrefetcher(dev, host_img[0], host_lab[0])

    def e2e_step(i):
        # the call a user makes: host batch in (pinned -> device), loss value out.  Every step's batch is
        # copied from pinned host memory inside the timed region -- batch 0 right here, batch i+1 on the copy
        # stream while step i computes (trainer.BatchPrefetcher, what DataLoader(pin_memory=True) + .to(device)
        # do in train.py).  The loss of every step is copied to pinned host memory and read on the host one
        # step later, so the device->host read does not drain the launch queue (the usual logging pattern).
        if i == 0:
            feeder.submit(host_img[0], host_lab[0])
        x, y = feeder.next()
        if i + 1 < args.steps:
            feeder.submit(host_img[(i + 1) % n_host], host_lab[(i + 1) % n_host])
        loss = step(x, y)
        k = i % 2
        loss_host[k].copy_(loss, non_blocking=True)
        loss_evt[k].record()
        if i > 0:
            loss_evt[1 - k].synchronize()
            e2e_losses.append(float(loss_host[1 - k]))

    # warm-up (also the first packs / allocator growth)
    _progress("host batches pinned and copied")
    for i in range(max(args.warmup, 3)):
        resident_step(i)
        if os.environ.get("KV_VERBOSE") == "1":
            torch.cuda.synchronize()
            _progress(f"warm-up step {i}")
    torch.cuda.synchronize()
    torch.cuda.reset_peak_memory_stats(dev)
    base_mem = torch.cuda.memory_allocated(dev)
    resident_step(0)
    torch.cuda.synchronize()
    step_ws_mb = (torch.cuda.max_memory_allocated(dev) - base_mem) / 1e6

    clocks = ClockSampler(local)
    clocks.start()
    l0 = lib.kv_launch_count()
    state["timed"] = True
    ms_total = timed(resident_step, args.steps)
    launches = lib.kv_launch_count() - l0
    _progress("resident steps timed")
    ms_e2e = timed(e2e_step, args.steps)
    _progress("e2e steps timed")
    e2e_losses.append(float(loss_host[(args.steps - 1) % 2]))  # the last step's loss (already on the host)
    clk = clocks.stop()

    # per-kernel shares for the roofline (separate pass: the events perturb the step slightly)
    l1 = lib.kv_launch_count()
    if args.cuda_graph:
        prof = profile_kernels(lambda: step._eager(dev_img[0], dev_lab[0]), 3)
        launches = (lib.kv_launch_count() - l1) // 3 * args.steps  # a replayed graph launches what it captured
    else:
        prof = profile_kernels(lambda: resident_step(0), 3)
    by_shape = prof.pop("_by_shape")
    peaks = load_peaks()
    kernel_ms = sum(v["ms_per_step"] for v in prof.values())
    top = max(prof, key=lambda k: prof[k]["ms_per_step"])
    t = prof[top]
    # dominant launch kind: entry point x layer shape for the GEMM-shaped kernels (their launches differ in
    # size by orders of magnitude), the entry point itself otherwise
    shaped = {k: v for k, v in by_shape.items() if k.startswith(top + " ")}
    if shaped:
        top_shape = max(shaped, key=lambda k: shaped[k]["ms_per_call"] * shaped[k]["calls"])
        ts = shaped[top_shape]
        top_name, ms_call, calls = top_shape, ts["ms_per_call"], ts["calls"]
        work_call = ts["tflops"] * 1e12 * ts["ms_per_call"] * 1e-3
    else:
        top_name, ms_call, calls, work_call = top, t["ms_per_call"], t["calls"], t["work_per_call"]
    if top in FLOP_KERNELS:
        achieved = work_call / (ms_call * 1e-3) / 1e12
        peak, unit, bound = peaks["bf16_tflops_sustained"], "TFLOP/s", "tensor"
    else:
        achieved = work_call / (ms_call * 1e-3) / 1e9
        peak, unit, bound = peaks["hbm_gbs"], "GB/s", "hbm"
    traffic = None
    for tname in ("r02_traffic.json", "r01_traffic.json"):
        tpath = os.path.join(ROOT, "profiles", tname)
        if os.path.exists(tpath):
            with open(tpath) as f:
                traffic = json.load(f).get(top_name)  # DRAM bytes of that launch, one ncu --set full capture
            break
    roofline = {"bound": bound, "achieved": achieved, "peak": peak, "unit": unit, "frac": achieved / peak,
                "traffic": traffic, "kernel": top_name, "kernel_ms_per_call": ms_call,
                "kernel_calls_per_step": calls, "kernel_share_of_step": ms_call * calls / kernel_ms,
                "algorithmic_work_per_call": work_call,
                "entry_point_mean": {"name": top, "achieved": t["work_per_call"] / (t["ms_per_call"] * 1e-3) /
                                     (1e12 if top in FLOP_KERNELS else 1e9), "share": t["ms_per_step"] / kernel_ms},
                "peak_source": peaks["source"] + " (sustained bf16 dense / copy bandwidth)",
                "shares": {k: round(v["ms_per_step"] / kernel_ms, 4) for k, v in
                           sorted(prof.items(), key=lambda kv: -kv[1]["ms_per_step"])},
                # the GEMM-shaped launches by layer shape (T rows, groups, in->out, M basis functions)
                "by_shape": {k: {"calls": v["calls"], "ms_per_call": round(v["ms_per_call"], 4),
                                 "tflops": round(v["tflops"], 1)}
                             for k, v in sorted(by_shape.items(), key=lambda kv: -kv[1]["ms_per_call"] * kv[1]["calls"])}}

    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cores = host_threads()
        n_cpu = 2 if name.startswith("cfg4") else 3
        cval, cms, n_done = time_cpu(name, cpu_batch, n_cpu, 1, 30.0)
        cpu_baseline = {"value": cval, "unit": UNIT, "cores": cores, "kind": "port",
                        "sample": f"{n_done} steps of batch {cpu_batch} of the same model and step "
                                  f"(oracle/kan_oracle.py, torch CPU fp32, {cms:.0f} ms/step)"}

    if rank == 0:
        ms_step = ms_total / args.steps
        img_b = host_img[0].numel() * host_img[0].element_size() + host_lab[0].numel() * host_lab[0].element_size()
        line = {
            "metric": METRIC, "value": world * batch / (ms_step * 1e-3), "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_step,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32" if args.precision == "fp32" else "bf16", "data": "synthetic",
            "config": {"workload": workload_name(name, batch), "global_batch": world * batch,
                       "parallelism": f"dp{world}", "cuda_graph": bool(args.cuda_graph),
                       "l2": f"no flush: one step touches {step_ws_mb:.0f} MB of activations > 126 MB L2"},
            "e2e": {"value": world * batch / (ms_e2e / args.steps * 1e-3), "unit": UNIT,
                    "h2d_bytes_per_step": img_b, "d2h_bytes_per_step": 4},
            "gpu_launches": int(launches),
            "clocks": clk,
            "roofline": roofline,
            "cpu_baseline": cpu_baseline,
        }
        print(json.dumps(line), file=_OUT, flush=True)
    if world > 1:
        dist.destroy_process_group()


# ---------------------------------------------------------------------------------------------
# --config cfg5: the "KANLinear fused TFLOP/s vs peak" half of BASELINE.json's metric.
# A step = forward + backward (input and coefficient gradients) of the two headline layers of configs[3]/[4]:
# KANLinear 768 -> 3072 and 3072 -> 768 (efficient-kan, grid 5, order 3, M = 9) on T = 65 536 token rows, chained the
# way the KAN MLP chains them, bf16 tensor-core arithmetic.  Algorithmic FLOPs 6 T in out M per layer.
# ---------------------------------------------------------------------------------------------
LAYER_T = 65536
LAYER_SHAPES = ((768, 3072), (3072, 768))
LAYER_M = 9
LAYER_CPU_T = 4096
LAYER_METRIC, LAYER_UNIT = "kan_linear_fwd_bwd_tflops", "TFLOP/s"


def layer_flops(T):
    return sum(6.0 * T * i * o * LAYER_M for i, o in LAYER_SHAPES)


def layer_workload(T):
    return (f"cfg5 headline layers: KANLinear 768->3072 -> KANLinear 3072->768 (efficient-kan grid 5 order 3), "
            f"T={T} token rows, step = forward + backward (dX and dW) of both layers")


def time_layer_cpu(T, steps, budget_s):
    """The reference algorithm of the two layers (oracle port: b_splines + two F.linear per layer, autograd backward)
    on the host cores.  TEST-INFRASTRUCTURE oracle used as the reported CPU baseline."""
    from oracle import kan_oracle as O
    g = torch.Generator().manual_seed(0)
    ws = []
    for i, o in LAYER_SHAPES:
        ws.append(((torch.randn(o, i, generator=g) * i ** -0.5).requires_grad_(),
                   (torch.randn(o, i, 8, generator=g) * 0.05).requires_grad_(),
                   torch.ones(o, i).requires_grad_(), O.effkan_default_grid(i)))
    x = torch.randn(T, LAYER_SHAPES[0][0], generator=g).requires_grad_()
    dy = torch.randn(T, LAYER_SHAPES[-1][1], generator=g)

    def step():
        h = x
        for bw, sw, sc, grid in ws:
            h = O.effkan_forward(h, bw, sw, sc, grid, 3)
        for t in [x] + [w for tup in ws for w in tup[:3]]:
            t.grad = None
        h.backward(dy)

    step()
    ts = []
    for _ in range(steps):
        t0 = time.perf_counter()
        step()
        ts.append(time.perf_counter() - t0)
        if sum(ts) > budget_s:
            break
    mean = sum(ts) / len(ts)
    return layer_flops(T) / mean / 1e12, mean * 1e3, len(ts)


SWEEP_SHAPES = ((192, 192), (384, 384), (768, 768), (1536, 1536), (3072, 3072), (768, 3072), (3072, 768))


def cpu_layer_sweep(args):
    """CPU column of the config-5 sweep (--impl reference --config cfg5 --sweep-cpu): forward + backward of ONE layer of
    each family (grid 5) on the host cores through the oracle port, T = 4096 rows (Fourier: 512, as the reference's
    own [T, out, in, G] intermediate would demand; the port uses the contraction form).  One JSON line per point."""
    from oracle import kan_oracle as O
    cores = host_threads()
    g = torch.Generator().manual_seed(0)
    rn = lambda *sh: (torch.randn(*sh, generator=g) * 0.05).requires_grad_()
    G = 5
    for fam in ("effkan", "cheby", "fastkan", "sine", "fourier"):
        for i, o in SWEEP_SHAPES:
            T = 512 if fam == "fourier" else 4096
            if fam == "effkan":
                ws = [rn(o, i), rn(o, i, G + 3), rn(o, i)]
                grid = O.effkan_default_grid(i, G, 3)
                f = lambda x: O.effkan_forward(x, ws[0], ws[1], ws[2], grid, 3)
                M = G + 4
            elif fam == "cheby":
                ws = [rn(i, o, G + 1)]
                f = lambda x: O.cheby_forward(x, ws[0])
                M = G + 1
            elif fam == "fastkan":
                ws = [torch.ones(i).requires_grad_(), torch.zeros(i).requires_grad_(), rn(o, i * G), rn(o, i), rn(o)]
                rg = torch.linspace(-2.0, 2.0, G)
                f = lambda x: O.fastkan_forward(x, ws[0], ws[1], rg, ws[2], ws[3], ws[4])
                M = G + 1
            elif fam == "sine":
                ws = [rn(o, i, G), O.sinekan_default_freq(G).requires_grad_(), rn(1, o)]
                ph = O.sinekan_phase(i, G)
                f = lambda x: O.sinekan_forward(x, ws[0], ws[1], ph, ws[2])
                M = G
            else:
                ws = [rn(2, o, i, G), rn(1, o)]
                f = lambda x: O.fourier_forward(x, ws[0], ws[1])
                M = 2 * G
            x = torch.randn(T, i, generator=g).requires_grad_()
            dy = torch.randn(T, o, generator=g)

            def step():
                for t in [x] + ws:
                    t.grad = None
                f(x).backward(dy)

            step()
            t0 = time.perf_counter()
            step()
            dt = time.perf_counter() - t0
            print(json.dumps({"family": fam, "T": T, "in": i, "out": o, "grid": G, "M": M, "impl": "cpu port (oracle)",
                              "cores": cores, "fwdbwd_ms": round(dt * 1e3, 2),
                              "fwdbwd_tflops": round(6.0 * T * i * o * M / dt / 1e12, 4)}), file=_OUT, flush=True)


def run_layer_reference_arm(args):
    if int(os.environ.get("RANK", "0")) != 0:
        return
    cores = host_threads()
    val, ms, done = time_layer_cpu(LAYER_CPU_T, args.steps, CPU_TIME_BUDGET_S)
    sample = (f"{done} steps at T={LAYER_CPU_T} rows (oracle/kan_oracle.py effkan_forward, torch CPU fp32, autograd; "
              f"{ms:.0f} ms/step)")
    line = {"metric": LAYER_METRIC, "value": val, "unit": LAYER_UNIT, "n_gpus": args.gpus, "steps": done, "warmup": 1,
            "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "impl": "reference",
            "config": {"workload": layer_workload(LAYER_T), "reference_sample": sample},
            "cpu_baseline": {"value": val, "unit": LAYER_UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": val, "unit": LAYER_UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), file=_OUT, flush=True)


def run_layer_arm(args):
    import torch.distributed as dist
    import kan_vit_b200 as K
    from kan_vit_b200 import _lib as L

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the GPU arm has no CPU fallback; use --impl reference)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)   # replicas only: the barrier and the max over ranks
    lib = L.load()
    K.set_precision("bf16")
    T = args.batch or LAYER_T
    torch.manual_seed(0)
    layers = [K.KANLinear(i, o).to(dev) for i, o in LAYER_SHAPES]
    g = torch.Generator().manual_seed(1000 + rank)
    n_host = 2
    host_x = [torch.randn(T, LAYER_SHAPES[0][0], generator=g).pin_memory() for _ in range(n_host)]
    host_dy = [torch.randn(T, LAYER_SHAPES[-1][1], generator=g).pin_memory() for _ in range(n_host)]
    dev_x = [t.to(dev).requires_grad_() for t in host_x]
    dev_dy = [t.to(dev) for t in host_dy]
    params = [p for m in layers for p in m.parameters()]

    def fwd_bwd(x, dy):
        for p in params:
            p.grad = None
        x.grad = None
        h = x
        for m in layers:
            h = m(h)
        h.backward(dy)
        return h

    def resident_step(i):
        fwd_bwd(dev_x[i % n_host], dev_dy[i % n_host])

    # end-to-end leg: rows and output gradient come from pinned host memory, the layer output and the input gradient go
    # back to pinned host memory, EVERY step, inside the timed region.  Copies run on two side streams against two
    # device slots (step i + 1's input lands while step i computes, step i's results leave while step i + 1 computes),
    # event-ordered, no host synchronisation; the last step's results are waited for before the closing event.
    main = torch.cuda.current_stream()
    s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()
    slot_x = [torch.empty_like(dev_x[0]).requires_grad_() for _ in range(2)]
    slot_dy = [torch.empty_like(dev_dy[0]) for _ in range(2)]
    ev_in = [torch.cuda.Event() for _ in range(2)]      # slot k holds its input
    ev_used = [torch.cuda.Event() for _ in range(2)]    # the step that read slot k has been issued and finished
    host_y = [torch.empty(T, LAYER_SHAPES[-1][1]).pin_memory() for _ in range(2)]
    host_dx = [torch.empty(T, LAYER_SHAPES[0][0]).pin_memory() for _ in range(2)]

    def submit(i):
        k = i % 2
        with torch.cuda.stream(s_in), torch.no_grad():
            if i >= 2:
                s_in.wait_event(ev_used[k])
            slot_x[k].copy_(host_x[i % n_host], non_blocking=True)
            slot_dy[k].copy_(host_dy[i % n_host], non_blocking=True)
            ev_in[k].record(s_in)

    def e2e_step(i):
        k = i % 2
        if i == 0:
            submit(0)
        if i + 1 < args.steps:
            submit(i + 1)
        main.wait_event(ev_in[k])
        y = fwd_bwd(slot_x[k], slot_dy[k])
        dx = slot_x[k].grad
        ev_used[k].record(main)
        with torch.cuda.stream(s_out):
            s_out.wait_event(ev_used[k])
            host_y[k].copy_(y.detach(), non_blocking=True)
            host_dx[k].copy_(dx, non_blocking=True)
            y.record_stream(s_out)
            dx.record_stream(s_out)
        if i + 1 == args.steps:
            main.wait_stream(s_out)      # the closing event sees the last results on the host

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            fn(i)
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    for i in range(max(args.warmup, 3)):
        resident_step(i)
    torch.cuda.synchronize()
    clocks = ClockSampler(local)
    clocks.start()
    l0 = lib.kv_launch_count()
    ms_total = timed(resident_step, args.steps)
    launches = lib.kv_launch_count() - l0
    ms_e2e = timed(e2e_step, args.steps)
    clk = clocks.stop()

    prof = profile_kernels(lambda: resident_step(0), 3)
    by_shape = prof.pop("_by_shape")
    peaks = load_peaks()
    gemm = {k: v for k, v in by_shape.items() if k.split(" ")[0] in FLOP_KERNELS}
    top = max(gemm, key=lambda k: gemm[k]["ms_per_call"] * gemm[k]["calls"])
    kernel_ms = sum(v["ms_per_step"] for v in prof.values())
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "r02_traffic.json")
    if os.path.exists(tpath):
        with open(tpath) as f:
            traffic = json.load(f).get(top.replace(f"T{T} ", "T66192 "))  # the same launch inside the cfg4 step
    roofline = {"bound": "tensor", "achieved": gemm[top]["tflops"], "peak": peaks["bf16_tflops_sustained"],
                "unit": "TFLOP/s", "frac": gemm[top]["tflops"] / peaks["bf16_tflops_sustained"], "traffic": traffic,
                "kernel": top, "kernel_ms_per_call": gemm[top]["ms_per_call"],
                "kernel_share_of_step": gemm[top]["ms_per_call"] * gemm[top]["calls"] / kernel_ms,
                "peak_source": peaks["source"] + " (sustained bf16 dense); burst " + str(peaks["bf16_tflops"]),
                "by_shape": {k: {"calls": v["calls"], "ms_per_call": round(v["ms_per_call"], 4), "tflops": round(v["tflops"], 1)}
                             for k, v in sorted(by_shape.items(), key=lambda kv: -kv[1]["ms_per_call"] * kv[1]["calls"])}}
    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cores = host_threads()
        cval, cms, n_done = time_layer_cpu(LAYER_CPU_T, 3, 30.0)
        cpu_baseline = {"value": cval, "unit": LAYER_UNIT, "cores": cores, "kind": "port",
                        "sample": f"{n_done} steps at T={LAYER_CPU_T} rows of the same two layers "
                                  f"(oracle/kan_oracle.py, torch CPU fp32, {cms:.0f} ms/step)"}
    if rank == 0:
        ms_step = ms_total / args.steps
        h2d = sum(t.numel() * t.element_size() for t in (host_x[0], host_dy[0]))
        d2h = sum(t.numel() * t.element_size() for t in (host_y[0], host_dx[0]))
        line = {"metric": LAYER_METRIC, "value": world * layer_flops(T) / (ms_step * 1e-3) / 1e12, "unit": LAYER_UNIT,
                "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_step,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
                "config": {"workload": layer_workload(T), "parallelism": f"replicas x{world}",
                           "l2": f"no flush: one step touches {4e-6 * T * (768 * 3 + 3072 * 3):.0f} MB of activations and "
                                 "gradients > 126 MB L2",
                           "frac_of_sustained_peak": (layer_flops(T) / (ms_step * 1e-3) / 1e12) / peaks["bf16_tflops_sustained"],
                           "frac_of_burst_peak": (layer_flops(T) / (ms_step * 1e-3) / 1e12) / peaks["bf16_tflops"]},
                "e2e": {"value": world * layer_flops(T) / (ms_e2e / args.steps * 1e-3) / 1e12, "unit": LAYER_UNIT,
                        "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
                "gpu_launches": int(launches), "clocks": clk, "roofline": roofline, "cpu_baseline": cpu_baseline}
        print(json.dumps(line), file=_OUT, flush=True)
    if world > 1:
        dist.destroy_process_group()


_OUT = sys.stdout


def main():
    # stdout carries exactly ONE JSON line: whatever libraries write to fd 1 (NCCL prints its version
    # banner there when NCCL_DEBUG is set on the box) is sent to stderr instead.
    global _OUT
    sys.stdout.flush()
    _OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="cfg4", choices=sorted(CONFIGS) + ["cfg5"])
    ap.add_argument("--batch", type=int, default=0, help="per-GPU batch (default: the config's)")
    ap.add_argument("--precision", default=None, choices=["fp32", "bf16"],
                    help="bf16 = tcgen05 / tensor-core kernels with fp32 accumulation, fp32 = exact SIMT kernels "
                         "(default: per config, bf16 for cfg3/cfg4)")
    ap.add_argument("--bf16", action="store_true", help="same as --precision bf16")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--cuda-graph", action="store_true",
                    help="single GPU: capture the whole iteration in a CUDA graph (launch-bound small configs)")
    ap.add_argument("--sweep-cpu", action="store_true",
                    help="with --config cfg5: the CPU column of the layer sweep (oracle port), one JSON line per point")
    args = ap.parse_args()
    if args.config == "cfg5" and args.sweep_cpu:
        if int(os.environ.get("RANK", "0")) == 0:
            cpu_layer_sweep(args)
        return
    if args.config == "cfg5":
        (run_layer_reference_arm if args.impl == "reference" else run_layer_arm)(args)
        return
    if args.precision is None:
        args.precision = "bf16" if args.bf16 else DEFAULT_PRECISION[args.config]
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_gpu_arm(args)


if __name__ == "__main__":
    main()
