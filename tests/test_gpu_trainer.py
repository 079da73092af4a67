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
