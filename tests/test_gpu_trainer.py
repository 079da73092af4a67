"""GPU checks of the training step plumbing: gradients written straight into the flat bucket buffer
(ops.GRAD_SINK) equal the ones autograd accumulates, and the CUDA-graph step equals the eager step."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _model(kind, dev):
    import kan_vit_b200 as K
    torch.manual_seed(0)
    return K.VisionTransformer((1, 28, 28), 7, 2, 64, 2, 10, kind, kan_mlp=(kind == "efficientkan")).to(dev)


@pytest.mark.parametrize("kind", ["efficientkan", "fast", "sine", "cheby", "vanilla"])
def test_grad_sink_matches_autograd(kind):
    from kan_vit_b200 import ops
    from kan_vit_b200.trainer import FlatGradBuckets
    dev = torch.device("cuda:0")
    m = _model(kind, dev)
    img = torch.randn(6, 1, 28, 28, device=dev)
    lab = torch.randint(0, 10, (6,), device=dev)
    buckets = FlatGradBuckets([p for p in m.parameters() if p.requires_grad])

    def run(use_sink):
        buckets.zero()
        loss = torch.nn.functional.cross_entropy(m(img), lab)
        ops.GRAD_SINK = buckets if use_sink else None
        try:
            loss.backward()
        finally:
            ops.GRAD_SINK = None
        return buckets.flat.clone()

    a, b = run(True), run(False)
    assert a.abs().max() > 0
    assert (a - b).abs().max().item() <= 1e-6 * b.abs().max().item()


def test_cuda_graph_step_matches_eager():
    from kan_vit_b200.trainer import TrainStep
    dev = torch.device("cuda:0")
    img = [torch.randn(8, 1, 28, 28, device=dev) for _ in range(6)]
    lab = [torch.randint(0, 10, (8,), device=dev) for _ in range(6)]
    losses = []
    for graph in (False, True):
        step = TrainStep(_model("efficientkan", dev), lr=1e-3, cuda_graph=graph)
        losses.append([float(step(x, y)) for x, y in zip(img, lab)])
    for a, b in zip(*losses):
        assert abs(a - b) <= 1e-4 * abs(a)


# ---- the loss and the optimizer kernels ----

@pytest.mark.parametrize("B,Cn,ignore", [(4, 10, False), (256, 10, False), (37, 100, True), (4096, 1000, True),
                                         (1, 3, False)])
def test_cross_entropy_matches_oracle_and_torch(B, Cn, ignore):
    import numpy as np
    from kan_vit_b200 import ops
    from oracle import kan_oracle as O
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(B + Cn)
    z = 4.0 * torch.randn(B, Cn, generator=g)
    y = torch.randint(0, Cn, (B,), generator=g)
    if ignore:
        y[::7] = -100
    loss, dz = ops.cross_entropy_with_grad(z.to(dev), y.to(dev))
    lo, dzo = O.cross_entropy_np(z.numpy(), y.numpy())
    assert abs(float(loss) - float(lo)) <= 2e-6 * abs(float(lo))          # fp32 tolerance, mean of B rows
    assert np.abs(dz.cpu().numpy() - dzo).max() <= 2e-7
    # torch on the same device, through the autograd wrapper
    zt = z.to(dev).requires_grad_(True)
    za = z.to(dev).requires_grad_(True)
    (3.0 * torch.nn.functional.cross_entropy(zt, y.to(dev))).backward()
    (3.0 * ops.cross_entropy(za, y.to(dev))).backward()
    assert (za.grad - zt.grad).abs().max().item() <= 1e-6
    # block partials are added in block order: bit-reproducible
    loss2, dz2 = ops.cross_entropy_with_grad(z.to(dev), y.to(dev))
    assert torch.equal(loss, loss2) and torch.equal(dz, dz2)


def test_cross_entropy_all_ignored_is_nan_and_bad_label_poisons():
    from kan_vit_b200 import ops
    dev = torch.device("cuda:0")
    z = torch.randn(5, 7, device=dev)
    loss, dz = ops.cross_entropy_with_grad(z, torch.full((5,), -100, device=dev))
    assert torch.isnan(loss) and dz.abs().max().item() == 0.0
    loss, _ = ops.cross_entropy_with_grad(z, torch.tensor([0, 1, 2, 9, 3], device=dev))
    assert torch.isnan(loss)
    loss, _ = ops.cross_entropy_with_grad(z, torch.tensor([0, 1, 2, 6, 3], device=dev))  # workspace left clean
    assert torch.isfinite(loss)


@pytest.mark.parametrize("n", [4, 1000, 1 << 20, 5 * (1 << 20) + 12])
def test_flat_adam_kernel_matches_oracle_and_torch(n):
    import numpy as np
    from kan_vit_b200 import ops
    from oracle import kan_oracle as O
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(n % 1000)
    p0 = torch.randn(n, generator=g)
    p = p0.to(dev).clone()
    m, v = torch.zeros(n, device=dev), torch.zeros(n, device=dev)
    state = torch.zeros(4, dtype=torch.float64, device=dev)
    pt = p0.to(dev).clone().requires_grad_(True)
    opt = torch.optim.Adam([pt], lr=3e-3)
    pn, mn, vn = p0.numpy().copy(), np.zeros(n, np.float32), np.zeros(n, np.float32)
    for t in range(1, 5):
        grad = torch.randn(n, generator=g) * (10.0 ** (t - 3))
        gd = (2.0 * grad).to(dev)                      # the kernel applies grad_scale = 1/2
        ops.adam_step(p, gd, m, v, state, 3e-3, grad_scale=0.5, zero_grad=(t % 2 == 0))
        assert (gd.abs().max().item() == 0.0) == (t % 2 == 0)
        pt.grad = grad.to(dev)
        opt.step()
        if n <= 1 << 20:
            pn, mn, vn = O.adam_step_np(pn, grad.numpy(), mn, vn, t, lr=3e-3)
            assert np.abs(p.cpu().numpy() - pn).max() <= 2e-6 * np.abs(pn).max()
        assert (p - pt.detach()).abs().max().item() <= 2e-6 * pt.detach().abs().max().item()
    assert float(state[0]) == 4.0


@pytest.mark.parametrize("kind", ["efficientkan", "sine"])
def test_fused_step_follows_torch_loss_and_adam(kind):
    """TrainStep with the library's cross-entropy + flat Adam kernels follows the same model trained with
    F.cross_entropy + torch.optim.Adam, and the parameters stay views of the flat buffer."""
    from kan_vit_b200.trainer import FlatAdam, TrainStep
    dev = torch.device("cuda:0")
    img = [torch.randn(8, 1, 28, 28, device=dev) for _ in range(4)]
    lab = [torch.randint(0, 10, (8,), device=dev) for _ in range(4)]
    ma, mb = _model(kind, dev), _model(kind, dev)
    fused = TrainStep(ma, lr=1e-3)
    assert isinstance(fused.opt, FlatAdam) and fused.fused
    plain = TrainStep(mb, optimizer=torch.optim.Adam(mb.parameters(), lr=1e-3))
    assert not plain.fused
    for x, y in zip(img, lab):
        la, lb = float(fused(x, y)), float(plain(x, y))
        assert abs(la - lb) <= 1e-4 * abs(lb)
    lo = fused.opt.flat_param.data_ptr()
    for (k, pa), pb in zip(ma.named_parameters(), mb.parameters()):
        assert lo <= pa.data_ptr() < lo + fused.opt.flat_param.numel() * 4, k
        # Adam's first steps move every weight by ~lr whatever the gradient scale, so compare against lr
        assert (pa - pb).abs().max().item() <= 2e-4, k
    assert fused.grads.flat.abs().max().item() == 0.0      # cleared by the Adam launch
    sd = ma.state_dict()
    mc = _model(kind, dev)
    mc.load_state_dict(sd)                                  # strict: names and shapes unchanged by the re-homing


# ---- metric collection and the input transform on the device ----

def test_cross_entropy_metric_outputs():
    from kan_vit_b200 import ops
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(9)
    B, Cn = 513, 100
    z = 3.0 * torch.randn(B, Cn, generator=g)
    z[5, 7] = z[5, 3] = z[5].max() + 1.0                  # a tie: torch.argmax takes the lowest index
    y = torch.randint(0, Cn, (B,), generator=g)
    probs = torch.empty(B, Cn, device=dev)
    pred = torch.empty(B, dtype=torch.int64, device=dev)
    conf = torch.zeros(Cn, Cn, dtype=torch.int32, device=dev)
    for _ in range(2):                                    # counts accumulate over steps
        ops.cross_entropy_with_grad(z.to(dev), y.to(dev), probs_out=probs, pred_out=pred, confusion=conf)
    assert (probs.cpu() - torch.softmax(z, 1)).abs().max().item() <= 1e-6
    assert torch.equal(pred.cpu(), torch.argmax(z, 1)) and pred[5].item() == 3
    ref = torch.zeros(Cn, Cn, dtype=torch.int32)
    ref.index_put_((y, torch.argmax(z, 1)), torch.ones(B, dtype=torch.int32), accumulate=True)
    assert torch.equal(conf.cpu(), 2 * ref)


def test_epoch_metrics_match_reference_recipe():
    """EpochMetrics over a short run equals utils.calculate_metrics' sklearn calls on the per-step host lists."""
    from sklearn.metrics import accuracy_score, balanced_accuracy_score, f1_score, roc_auc_score
    from kan_vit_b200.metrics import EpochMetrics
    from kan_vit_b200.trainer import TrainStep
    dev = torch.device("cuda:0")
    m = _model("vanilla", dev)
    em = EpochMetrics(10, 64, dev)
    step = TrainStep(m, lr=1e-3, metrics=em)
    y_true, y_pred, y_proba, losses = [], [], [], []
    g = torch.Generator().manual_seed(2)
    for _ in range(4):
        x = torch.randn(16, 1, 28, 28, generator=g).to(dev)
        y = torch.randint(0, 10, (16,), generator=g).to(dev)
        with torch.no_grad():
            y_hat = m(x)                                   # what the step is about to see (train.py:33)
        losses.append(float(step(x, y)))
        y_true.extend(y.cpu().numpy())
        y_pred.extend(torch.argmax(y_hat, dim=1).cpu().numpy())
        y_proba.extend(torch.softmax(y_hat, dim=1).cpu().numpy())
    out = em.compute()
    assert abs(out["loss"] - sum(losses) / 4) < 1e-6
    assert abs(out["accuracy"] - accuracy_score(y_true, y_pred)) < 1e-12
    assert abs(out["balanced_accuracy"] - balanced_accuracy_score(y_true, y_pred)) < 1e-12
    assert abs(out["f1"] - f1_score(y_true, y_pred, average="weighted")) < 1e-12
    onehot = torch.nn.functional.one_hot(torch.tensor(y_true), num_classes=10).numpy()
    assert abs(out["roc_auc"] - roc_auc_score(onehot, y_proba, average="weighted", multi_class="ovr")) < 1e-6


@pytest.mark.parametrize("train", [True, False])
def test_image_augment_matches_torchvision(train):
    """train.py:99-109: RandomHorizontalFlip -> RandomCrop(32, padding=4) -> ToTensor -> Normalize, with the
    random draws fixed; bit-for-bit the torchvision arithmetic up to one rounding of the division."""
    tvf = pytest.importorskip("torchvision.transforms.functional")
    from kan_vit_b200 import ops
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(5)
    B, H, W, Cc, pad = 9, 32, 32, 3, 4
    img = torch.randint(0, 256, (B, H, W, Cc), dtype=torch.uint8, generator=g)
    mean, std = [0.5071, 0.4867, 0.4408], [0.2675, 0.2565, 0.2761]
    flip = torch.rand(B, generator=g) < 0.5
    top = torch.randint(0, 2 * pad + 1, (B,), generator=g, dtype=torch.int32)
    left = torch.randint(0, 2 * pad + 1, (B,), generator=g, dtype=torch.int32)
    top[0], left[0], top[1], left[1] = 0, 0, 2 * pad, 2 * pad
    ref = []
    for b in range(B):
        t = img[b].permute(2, 0, 1)                         # uint8 CHW, what PIL -> the transforms see
        if train:
            if flip[b]:
                t = tvf.hflip(t)
            t = tvf.crop(tvf.pad(t, pad), int(top[b]), int(left[b]), H, W)
        ref.append(tvf.normalize(t.float().div(255), mean, std))
    ref = torch.stack(ref)
    md, sd = torch.tensor(mean, device=dev), torch.tensor(std, device=dev)
    if train:
        out = ops.image_augment(img.to(dev), md, sd, pad=pad, flip=flip.to(dev), top=top.to(dev), left=left.to(dev))
    else:
        out = ops.image_augment(img.to(dev), md, sd)
    assert out.shape == (B, Cc, H, W)
    assert (out.cpu() - ref).abs().max().item() <= 1e-6


def test_batch_prefetcher_feeds_the_same_batches():
    """Batches copied one step ahead on the copy stream arrive intact and in order, slots are not overwritten
    before the step that reads them has run, and training through the prefetcher equals training on resident data."""
    from kan_vit_b200.trainer import BatchPrefetcher, TrainStep
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(11)
    imgs = [torch.randn(8, 1, 28, 28, generator=g).pin_memory() for _ in range(5)]
    labs = [torch.randint(0, 10, (8,), generator=g).pin_memory() for _ in range(5)]
    losses = []
    for prefetch in (False, True):
        step = TrainStep(_model("efficientkan", dev), lr=1e-3)
        out = []
        if prefetch:
            feeder = BatchPrefetcher(dev, imgs[0], labs[0])
            feeder.submit(imgs[0], labs[0])
            for i in range(5):
                x, y = feeder.next()
                if i + 1 < 5:
                    feeder.submit(imgs[i + 1], labs[i + 1])
                assert torch.equal(x.cpu(), imgs[i]) and torch.equal(y.cpu(), labs[i])
                out.append(float(step(x, y)))
            with pytest.raises(RuntimeError):
                feeder.next()
        else:
            for x, y in zip(imgs, labs):
                out.append(float(step(x.to(dev), y.to(dev))))
        losses.append(out)
    for a, b in zip(*losses):      # atomically accumulated gradient sums are not bit-reproducible run to run
        assert abs(a - b) <= 1e-5 * abs(a)


# ---- the fused multi-GPU step (gradient sink -> NCCL bucket all-reduces -> one Adam launch): 2 GPUs ----
def _nccl_worker(rank, world, port, out_dir):
    import os
    import sys
    import torch.distributed as dist
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    import kan_vit_b200 as K
    from kan_vit_b200.trainer import TrainStep
    torch.manual_seed(50 + rank)    # different seeds on purpose: TrainStep must broadcast rank 0's replica
    m = K.VisionTransformer((1, 28, 28), 7, 2, 64, 2, 10, "efficientkan", kan_mlp=True).to(dev)
    step = TrainStep(m, lr=1e-3, bucket_bytes=64 << 10)
    assert step.fused and len(step.grads.buckets) > 2
    g = torch.Generator().manual_seed(3)
    img = torch.randn(8, 1, 28, 28, generator=g)
    lab = torch.randint(0, 10, (8,), generator=g)
    xs, ys = img.chunk(world)[rank].to(dev), lab.chunk(world)[rank].to(dev)
    losses = [float(step(xs, ys)) for _ in range(3)]
    torch.save({"sd": {k: v.cpu() for k, v in m.state_dict().items()}, "losses": losses},
               os.path.join(out_dir, f"n{rank}.pt"))
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs (gpurun --gpus 2)")
def test_fused_two_gpu_step_matches_single_process(tmp_path):
    import os
    import socket
    import torch.multiprocessing as mp
    import kan_vit_b200 as K
    from kan_vit_b200.trainer import TrainStep
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    mp.spawn(_nccl_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    dev = torch.device("cuda:0")
    torch.manual_seed(50)           # rank 0's replica
    m = K.VisionTransformer((1, 28, 28), 7, 2, 64, 2, 10, "efficientkan", kan_mlp=True).to(dev)
    step = TrainStep(m, lr=1e-3)
    g = torch.Generator().manual_seed(3)
    img = torch.randn(8, 1, 28, 28, generator=g).to(dev)
    lab = torch.randint(0, 10, (8,), generator=g).to(dev)
    for _ in range(3):
        step(img, lab)
    r0, r1 = (torch.load(os.path.join(tmp_path, f"n{r}.pt")) for r in range(2))
    for k, v in m.state_dict().items():
        assert torch.equal(r0["sd"][k], r1["sd"][k]), k                    # replicas stay identical
        if v.is_floating_point():
            # Adam's first steps move every weight by ~lr whatever the gradient scale, so fp32 summation-order
            # differences in tiny gradients show up at the 1e-5 level of a 1e-3 update
            assert (r0["sd"][k] - v.cpu()).abs().max().item() <= 2e-5 + 1e-4 * v.abs().max().item(), k
