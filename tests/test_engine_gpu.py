"""TrainStep on one GPU: the captured CUDA graph (weight gradients on a second branch of the
graph, see ops/functional.py wgrad_stream) must train exactly like eager execution."""
import copy

import pytest
import torch

pytestmark = pytest.mark.gpu


def _train(model, graph, steps, shape, seed):
    import distributed_training_dl_b200 as dtdl
    from distributed_training_dl_b200.engine import CrossEntropyLoss, TrainStep
    net = dtdl.wrap(model.cuda())
    opt = dtdl.FusedSGD(net.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-4)
    step = TrainStep(net, opt, CrossEntropyLoss(), graph=graph, warmup=2)
    g = torch.Generator().manual_seed(seed)
    losses = []
    for _ in range(steps):
        x = torch.randn(*shape, generator=g).pin_memory()
        t = torch.randint(0, 10, (shape[0],), generator=g).pin_memory()
        losses.append(step(x, t).item())
    torch.cuda.synchronize()
    return losses, [p.detach().float().cpu().clone() for p in net.parameters()], step


def test_graph_step_equals_eager_step_without_batchnorm():
    """MnistCNN: the only float atomics are the loss sum and the bias-gradient column sums, so 8
    steps through the graph -- 6 of them replays with forked weight-gradient kernels -- agree
    with 8 eager steps to the last few bits."""
    from distributed_training_dl_b200.models import MnistCNN
    torch.manual_seed(0)
    base = MnistCNN()
    le, pe, _ = _train(copy.deepcopy(base), False, 8, (32, 1, 28, 28), 3)
    lg, pg, step = _train(copy.deepcopy(base), True, 8, (32, 1, 28, 28), 3)
    assert step.launches_per_step and step.launches_per_step > 0      # the graph was captured
    assert all(abs(a - b) < 1e-5 for a, b in zip(le, lg)), (le, lg)
    for a, b in zip(pe, pg):
        assert torch.allclose(a, b, rtol=1e-4, atol=1e-6), (a - b).abs().max()


def test_graph_step_tracks_eager_step_with_batchnorm():
    """With BatchNorm the statistics are summed with float atomics, so runs differ in the last
    bits; the loss curves of graph and eager training must still agree closely."""
    from distributed_training_dl_b200.models import PyramidNet
    torch.manual_seed(0)
    base = PyramidNet(num_layers=3, alpha=45.0)
    le, _, _ = _train(copy.deepcopy(base), False, 8, (16, 3, 32, 32), 5)
    lg, _, _ = _train(copy.deepcopy(base), True, 8, (16, 3, 32, 32), 5)
    assert all(l == l for l in lg)
    for a, b in zip(le, lg):
        assert abs(a - b) < 0.05 * max(1.0, abs(a)), (le, lg)
