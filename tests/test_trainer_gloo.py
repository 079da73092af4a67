"""Host-side data-parallel logic (kan_vit_b200/trainer.py) on CPU: world_size 2 over gloo.

The flat-bucket gradient all-reduce must give every rank the mean of the per-rank gradients, and
two ranks training on two half-batches must follow the single-process full-batch trajectory."""
import os
import socket
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _model():
    torch.manual_seed(0)
    return torch.nn.Sequential(torch.nn.Linear(12, 32), torch.nn.Tanh(), torch.nn.Linear(32, 32), torch.nn.Tanh(),
                               torch.nn.Linear(32, 5))


def _data():
    g = torch.Generator().manual_seed(1)
    return torch.randn(16, 12, generator=g), torch.randint(0, 5, (16,), generator=g)


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from kan_vit_b200.trainer import TrainStep
    m = _model()
    # tiny buckets so several collectives are issued during the backward pass
    step = TrainStep(m, lr=1e-2, bucket_bytes=1024, optimizer=torch.optim.SGD(m.parameters(), lr=1e-2))
    assert len(step.grads.buckets) > 1
    x, y = _data()
    xs, ys = x.chunk(world)[rank], y.chunk(world)[rank]
    for _ in range(3):
        step(xs, ys)
    torch.save({k: v.clone() for k, v in m.state_dict().items()}, os.path.join(out_dir, f"r{rank}.pt"))
    dist.destroy_process_group()


def test_two_ranks_match_single_process(tmp_path):
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    sys.path.insert(0, ROOT)
    from kan_vit_b200.trainer import TrainStep
    m = _model()
    step = TrainStep(m, optimizer=torch.optim.SGD(m.parameters(), lr=1e-2))
    x, y = _data()
    for _ in range(3):
        step(x, y)
    r0 = torch.load(os.path.join(tmp_path, "r0.pt"))
    r1 = torch.load(os.path.join(tmp_path, "r1.pt"))
    for k, v in m.state_dict().items():
        assert torch.equal(r0[k], r1[k]), k                     # replicas stay identical
        assert torch.allclose(r0[k], v, rtol=1e-5, atol=1e-6), k  # mean of half-batch grads = full-batch grad


def test_flat_views_and_zero():
    sys.path.insert(0, ROOT)
    from kan_vit_b200.trainer import FlatGradBuckets
    m = _model()
    fb = FlatGradBuckets(list(m.parameters()), bucket_bytes=256)
    x, y = _data()
    fb.zero()
    torch.nn.functional.cross_entropy(m(x), y).backward()
    for p in m.parameters():
        assert p.grad.data_ptr() >= fb.flat.data_ptr()
        assert p.grad.data_ptr() < fb.flat.data_ptr() + fb.flat.numel() * 4
    assert fb.flat.abs().sum() > 0
    fb.zero()
    assert all(float(p.grad.abs().sum()) == 0.0 for p in m.parameters())


# ---- a tied parameter reports in once per use; its bucket is reduced after the last one ----
class _SinkTwice(torch.autograd.Function):
    """Stands in for a KAN layer under ops.GRAD_SINK: writes its weight gradient straight into the flat bucket
    view and calls done(), instead of returning it to autograd."""

    @staticmethod
    def forward(ctx, x, w, sink):
        ctx.save_for_backward(x, w)
        ctx.sink = sink
        return x @ w.t()

    @staticmethod
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        ctx.sink.target(w).add_(dy.t() @ x)
        ctx.sink.done(w)
        return dy @ w, None, None


def _tied_worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from kan_vit_b200.trainer import FlatGradBuckets
    torch.manual_seed(0)
    w = torch.nn.Parameter(torch.randn(6, 6) * 0.3)
    x, _ = _data()
    xs = x[:, :6].chunk(world)[rank]
    fb = FlatGradBuckets([w], bucket_bytes=64)
    flats = []
    for _ in range(3):   # iteration 0 learns the use count, 1 and 2 fire the bucket from done()
        fb.zero()
        h = _SinkTwice.apply(xs, w, fb)
        _SinkTwice.apply(torch.tanh(h), w, fb).sum().backward()
        fb.finish()
        flats.append(fb.flat.clone())
    # two done() calls, plus autograd's post-accumulate hook, which fires once even though no gradient is returned
    assert fb._expected_uses[id(w)] >= 2, (fb._expected_uses, fb._uses)
    torch.save(flats, os.path.join(out_dir, f"t{rank}.pt"))
    dist.destroy_process_group()


def test_tied_parameter_reduced_after_last_use(tmp_path):
    port = _free_port()
    mp.spawn(_tied_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    torch.manual_seed(0)
    w = torch.randn(6, 6) * 0.3
    w.requires_grad_(True)
    x, _ = _data()
    (torch.tanh(x[:, :6] @ w.t()) @ w.t()).sum().backward()
    want = (w.grad / 2).reshape(-1)   # mean over the two ranks of the per-shard sums
    for r in range(2):
        for got in torch.load(os.path.join(tmp_path, f"t{r}.pt")):
            assert torch.allclose(got[:36], want, rtol=1e-5, atol=1e-6)


def _bcast_worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from kan_vit_b200.trainer import TrainStep
    torch.manual_seed(100 + rank)        # replicas built under DIFFERENT seeds
    m = torch.nn.Linear(4, 3)
    TrainStep(m, optimizer=torch.optim.SGD(m.parameters(), lr=1e-2))
    torch.save(m.state_dict(), os.path.join(out_dir, f"b{rank}.pt"))
    dist.destroy_process_group()


def test_trainstep_broadcasts_rank0_parameters(tmp_path):
    port = _free_port()
    mp.spawn(_bcast_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    a, b = (torch.load(os.path.join(tmp_path, f"b{r}.pt")) for r in range(2))
    torch.manual_seed(100)
    want = torch.nn.Linear(4, 3).state_dict()
    for k in want:
        assert torch.equal(a[k], want[k]) and torch.equal(b[k], want[k]), k


def test_cuda_graph_argument_errors():
    import pytest
    sys.path.insert(0, ROOT)
    from kan_vit_b200.trainer import TrainStep
    with pytest.raises(ValueError):
        TrainStep(_model(), cuda_graph=True)          # CPU model
