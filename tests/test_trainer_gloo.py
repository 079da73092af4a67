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
