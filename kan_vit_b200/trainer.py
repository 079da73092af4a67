"""Data-parallel training step around the hot path (the step shape of train.py:31-40 of the
reference: zero_grad -> forward -> CrossEntropyLoss -> backward -> Adam.step), one process per GPU.

The reference has no multi-GPU code at all; the path shards over the batch,
so each rank holds a full replica and its own slice of the batch, and the only exchange step per
iteration is the gradient all-reduce.  Gradients live in ONE flat fp32 buffer (every ``p.grad`` is
a view into it), cut into buckets in reverse parameter order; a bucket's all-reduce is issued on
the collective stream as soon as autograd has produced its last gradient, so the transfer
over NVLink hides under the rest of the backward pass.  Works with any ``torch.distributed``
backend (NCCL on the GPUs, gloo in the CPU tests).
"""
from __future__ import annotations

from typing import List, Optional

import torch
import torch.distributed as dist
import torch.nn.functional as F


class FlatGradBuckets:
    """All gradients of ``params`` as views into one flat buffer, all-reduced bucket by bucket."""

    def __init__(self, params: List[torch.nn.Parameter], bucket_bytes: int = 32 << 20, group=None):
        self.params = [p for p in params if p.requires_grad]
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
        dev, dt = self.params[0].device, self.params[0].dtype
        # reverse order: autograd produces the gradients of the last layers first
        order = list(reversed(self.params))
        offs, total = [], 0
        for p in order:
            offs.append(total)
            total += (p.numel() + 3) // 4 * 4
        self.flat = torch.zeros(total, dtype=dt, device=dev)
        self.order, self.offsets = order, offs
        self.buckets = []          # (start, end, n_params)
        self._bucket_of = {}
        start, count, elems_per_bucket = 0, 0, max(bucket_bytes // self.flat.element_size(), 1)
        for p, off in zip(order, offs):
            p.grad = self.flat[off:off + p.numel()].view_as(p)
            end = off + (p.numel() + 3) // 4 * 4
            self._bucket_of[id(p)] = len(self.buckets)
            count += 1
            if end - start >= elems_per_bucket:
                self.buckets.append((start, end, count))
                start, count = end, 0
        if count:
            self.buckets.append((start, total, count))
        self._pending = [0] * len(self.buckets)
        self._handles = []
        self._hooks = []
        # Events per parameter per iteration.  A parameter reports in through ops.GRAD_SINK `done` once per use (a
        # tied parameter applied by two layers: twice) AND through autograd's post-accumulate hook, which fires
        # once per backward pass even when a layer returned no gradient for it (it wrote into the bucket view
        # itself) -- always after the last use.  The bucket may only be reduced after the LAST event, so the
        # counts are learnt in the first iteration (no early all-reduce there: finish() reduces everything)
        # and checked afterwards.
        self._uses = {}
        self._expected_uses = None
        if self.world > 1:
            for p in self.params:
                self._hooks.append(p.register_post_accumulate_grad_hook(self._on_grad))

    def zero(self, already_zero: bool = False):
        """Start an iteration.  ``already_zero``: the optimizer kernel cleared the buffer as it read it."""
        if not already_zero:
            self.flat.zero_()
        for i, (_, _, n) in enumerate(self.buckets):
            self._pending[i] = n
        self._handles = []
        self._uses = {}

    def _on_grad(self, p):
        b = self._bucket_of[id(p)]
        n = self._uses[id(p)] = self._uses.get(id(p), 0) + 1
        if self._expected_uses is None:
            return                      # first iteration: learn the use counts, reduce in finish()
        want = self._expected_uses.get(id(p), 1)
        if n < want:
            return
        if n > want:
            raise RuntimeError("a parameter received more gradient contributions than in the previous iteration; "
                               "its bucket was already all-reduced (rebuild TrainStep after changing the graph)")
        self._pending[b] -= 1
        if self._pending[b] == 0:
            s, e, _ = self.buckets[b]
            self._handles.append(dist.all_reduce(self.flat[s:e], op=dist.ReduceOp.SUM, group=self.group,
                                                 async_op=True))

    # -- ops.GRAD_SINK protocol: the KAN layers add their parameter gradients straight into the flat buffer
    def target(self, p):
        return p.grad if id(p) in self._bucket_of else None

    def done(self, p):
        if self.world > 1:
            self._on_grad(p)

    def finish(self, scale: bool = True):
        """Wait for the collectives, reduce buckets whose hooks never fired (unused parameters),
        and turn the sums into means (``scale=False``: the optimizer kernel applies the 1/world)."""
        if self.world == 1:
            return
        if self._expected_uses is None or any(self._uses.get(k, 0) < v for k, v in self._expected_uses.items()):
            self._expected_uses = {k: max(v, 1) for k, v in self._uses.items()}
        for b, left in enumerate(self._pending):
            if left > 0:
                s, e, _ = self.buckets[b]
                self._handles.append(dist.all_reduce(self.flat[s:e], op=dist.ReduceOp.SUM,
                                                     group=self.group, async_op=True))
                self._pending[b] = 0
        for h in self._handles:
            h.wait()
        self._handles = []
        if scale:
            self.flat.mul_(1.0 / self.world)


class FlatAdam:
    """``torch.optim.Adam(params, lr)`` (train.py:23) as ONE kernel launch per iteration (kv_adam_step).

    The parameters are re-homed into one flat buffer laid out like the gradient buffer of ``grads``
    (every ``p.data`` becomes a view, values kept), the two moments are flat buffers of the same
    layout, and the step counter lives on the device.  The launch also applies the 1/world of the
    data-parallel mean and clears the gradient buffer for the next iteration.  CUDA only."""

    def __init__(self, grads: FlatGradBuckets, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8):
        if not grads.flat.is_cuda:
            raise RuntimeError("FlatAdam runs on CUDA buffers only (there is no CPU path)")
        self.grads = grads
        self.lr, self.betas, self.eps = float(lr), (float(betas[0]), float(betas[1])), float(eps)
        self.flat_param = torch.zeros_like(grads.flat)
        with torch.no_grad():
            for p, off in zip(grads.order, grads.offsets):
                view = self.flat_param[off:off + p.numel()].view_as(p)
                view.copy_(p.data)
                p.data = view
        self.exp_avg = torch.zeros_like(grads.flat)
        self.exp_avg_sq = torch.zeros_like(grads.flat)
        self.state = torch.zeros(4, dtype=torch.float64, device=grads.flat.device)

    def step(self, zero_grad: bool = True):
        from . import ops
        ops.adam_step(self.flat_param, self.grads.flat, self.exp_avg, self.exp_avg_sq, self.state, self.lr,
                      self.betas[0], self.betas[1], self.eps, grad_scale=1.0 / self.grads.world,
                      zero_grad=zero_grad)

    def state_dict(self):
        return {"exp_avg": self.exp_avg, "exp_avg_sq": self.exp_avg_sq, "state": self.state,
                "lr": self.lr, "betas": self.betas, "eps": self.eps}

    def load_state_dict(self, sd):
        self.exp_avg.copy_(sd["exp_avg"])
        self.exp_avg_sq.copy_(sd["exp_avg_sq"])
        self.state.copy_(sd["state"])
        self.lr, self.betas, self.eps = float(sd["lr"]), tuple(sd["betas"]), float(sd["eps"])


class BatchPrefetcher:
    """Host batches to the device one step ahead of the compute (the job of the reference's
    ``DataLoader(pin_memory=True)`` + ``x.to(device)``, train.py:31-32,115-116).

    Two static device slots; ``submit`` copies a pinned host batch into the free slot on a copy stream,
    ``next`` makes the compute stream wait for the oldest submitted batch and returns its device tensors.
    The copy of batch i+1 therefore runs under the kernels of step i.  Stream-ordered only: no host sync."""

    def __init__(self, device, images_like: torch.Tensor, labels_like: torch.Tensor):
        self.device = torch.device(device)
        self.stream = torch.cuda.Stream(self.device)
        self.img = [torch.empty(images_like.shape, dtype=images_like.dtype, device=self.device) for _ in range(2)]
        self.lab = [torch.empty(labels_like.shape, dtype=labels_like.dtype, device=self.device) for _ in range(2)]
        self.ready = [torch.cuda.Event(), torch.cuda.Event()]
        self.consumed = [None, None]     # recorded on the compute stream once the NEXT batch is taken
        self.head = self.tail = 0        # batches submitted / taken

    def submit(self, images: torch.Tensor, labels: torch.Tensor):
        if self.head - self.tail >= 2:
            raise RuntimeError("both slots hold batches that have not been taken yet")
        k = self.head % 2
        if self.consumed[k] is not None:
            self.stream.wait_event(self.consumed[k])   # the step that read this slot has finished
        with torch.cuda.stream(self.stream):
            self.img[k].copy_(images, non_blocking=True)
            self.lab[k].copy_(labels, non_blocking=True)
            self.ready[k].record(self.stream)
        self.head += 1

    def next(self):
        if self.tail >= self.head:
            raise RuntimeError("no batch submitted")
        cur = torch.cuda.current_stream(self.device)
        if self.tail > 0:
            # everything queued so far on the compute stream (the previous step) read the previous slot
            prev = (self.tail - 1) % 2
            self.consumed[prev] = torch.cuda.Event()
            self.consumed[prev].record(cur)
        k = self.tail % 2
        cur.wait_event(self.ready[k])
        self.tail += 1
        return self.img[k], self.lab[k]


class TrainStep:
    """One training iteration of a model on this rank's batch shard."""

    def __init__(self, model: torch.nn.Module, lr: float = 1e-3, bucket_bytes: int = 32 << 20,
                 optimizer: Optional[torch.optim.Optimizer] = None, cuda_graph: bool = False, metrics=None):
        """``cuda_graph=True`` (single-GPU, CUDA only): after three eager warm-up calls the whole
        iteration (zero_grad .. Adam) is captured once into a CUDA graph and replayed, which removes
        the launch-bound host time of the small configurations (thousands of launches of a few
        microseconds each).  Inputs are copied into static buffers; shapes must not change."""
        self.model = model
        # metrics: a kan_vit_b200.metrics.EpochMetrics; the loss launch then also records the batch's
        # probabilities / predictions / confusion counts on the device (train.py:42-44)
        self.metrics = metrics
        if metrics is not None and cuda_graph:
            raise ValueError("metric collection advances host-side offsets every step and cannot be graph-captured")
        params = [p for p in model.parameters() if p.requires_grad]
        self.grads = FlatGradBuckets(params, bucket_bytes)
        if cuda_graph and not params[0].is_cuda:
            raise ValueError("cuda_graph=True needs a CUDA model")
        if cuda_graph and optimizer is not None:
            raise ValueError("cuda_graph=True captures the library's own Adam launch (device-side step count); a "
                             "caller-supplied torch optimizer is not capturable here")
        if cuda_graph and self.grads.world > 1:
            raise ValueError("cuda_graph=True is single-GPU: the bucketed all-reduces are issued from host-side "
                             "bookkeeping during backward and are not captured")
        self.cuda_graph = bool(cuda_graph)
        # On the GPU the loss and the update are the library's own kernels: one
        # cross-entropy launch (loss + d logits) and one Adam launch over the flat buffers.  A caller's
        # optimizer, or CPU tensors (the gloo tests of the host logic), take the torch route.
        self.fused = optimizer is None and params[0].is_cuda
        if self.fused:
            optimizer = FlatAdam(self.grads, lr=lr)
        elif optimizer is None:
            optimizer = torch.optim.Adam(params, lr=lr)
        self.opt = optimizer
        if self.grads.world > 1:
            # replicas must start identical: rank 0's parameters win (models built under different seeds would
            # otherwise diverge silently)
            with torch.no_grad():
                if self.fused:
                    dist.broadcast(self.opt.flat_param, 0, group=self.grads.group)
                else:
                    for p in params:
                        dist.broadcast(p.data, 0, group=self.grads.group)
                for b in model.buffers():
                    if b.is_floating_point():
                        dist.broadcast(b, 0, group=self.grads.group)
        self._calls = 0
        self._clean = True         # gradient buffer known to be all zero
        self._graph = None
        self._static = None

    def _eager(self, images: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        # fused: the previous Adam launch left the gradient buffer cleared (it is created zeroed)
        self.grads.zero(already_zero=self.fused and self._clean)
        self._clean = False
        logits = self.model(images)
        if images.is_cuda:
            from . import ops
            if self.metrics is not None:
                loss, dlogits = self.metrics.loss_and_grad(logits.detach(), labels)
            else:
                loss, dlogits = ops.cross_entropy_with_grad(logits.detach(), labels)
            ops.GRAD_SINK = self.grads
            try:
                logits.backward(dlogits)
            finally:
                ops.GRAD_SINK = None
        else:
            loss = F.cross_entropy(logits, labels)
            loss.backward()
        self.grads.finish(scale=not self.fused)
        self.opt.step()
        self._clean = True
        return loss.detach()

    def __call__(self, images: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        if not self.cuda_graph:
            return self._eager(images, labels)
        self._calls += 1
        if self._calls <= 3:  # warm-up: allocator growth, pointer tables, grid checks (host syncs)
            return self._eager(images, labels)
        if self._graph is None:
            self._static = (images.clone(), labels.clone())
            torch.cuda.synchronize()
            self._graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self._graph):
                self._static_loss = self._eager(*self._static)
        self._static[0].copy_(images, non_blocking=True)
        self._static[1].copy_(labels, non_blocking=True)
        self._graph.replay()
        return self._static_loss
