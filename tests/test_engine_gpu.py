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
    # two runs of the same configuration already differ in the last bits (float atomics), and eight
    # SGD steps amplify that: a wrong kernel shows up as NaN or an O(1) gap, not as a few percent
    for a, b in zip(le, lg):
        assert abs(a - b) < 0.10 * max(1.0, abs(a)), (le, lg)


@pytest.mark.parametrize("name", ["mnist_cnn", "pyramidnet_shallow"])
def test_loss_curve_tracks_stock_torch_training(name):
    """N training steps of this framework (bf16 kernels, fused SGD, CUDA
    graph) against the same model trained by stock PyTorch in fp32 on the same data and seed."""
    from distributed_training_dl_b200.models import MnistCNN, PyramidNet
    torch.manual_seed(0)
    if name == "mnist_cnn":
        base, shape, tol = MnistCNN(), (32, 1, 28, 28), 0.03
    else:
        base, shape, tol = PyramidNet(num_layers=3, alpha=45.0), (16, 3, 32, 32), 0.15
    steps = 10
    ours, _, _ = _train(copy.deepcopy(base), True, steps, shape, 7)

    ref = copy.deepcopy(base).cuda()          # non-native inputs take the stock torch path
    opt = torch.optim.SGD(ref.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-4)
    g = torch.Generator().manual_seed(7)
    theirs = []
    for _ in range(steps):
        x = torch.randn(*shape, generator=g)
        t = torch.randint(0, 10, (shape[0],), generator=g)
        with torch.backends.cudnn.flags(allow_tf32=False):
            loss = torch.nn.functional.cross_entropy(_stock_forward(ref, x.cuda()), t.cuda())
        opt.zero_grad()
        loss.backward()
        opt.step()
        theirs.append(loss.item())
    assert all(v == v for v in theirs)
    for a, b in zip(ours, theirs):
        assert abs(a - b) < tol * max(1.0, abs(b)), (ours, theirs)


def _stock_forward(model, x):
    """Runs a framework model through its plain-PyTorch branch on the GPU (the branch the CPU
    oracle uses), by marking the input as not native."""
    from distributed_training_dl_b200.ops import modules as L
    orig = L.is_native
    L.is_native = lambda t: False
    try:
        return model(x)
    finally:
        L.is_native = orig


def test_pinned_prefetcher_feeds_async_copies():
    """Page-locked ring + event hand-back: 40 batches copied to the device with non_blocking copies
    while the worker keeps refilling the same three buffers -- every device batch must be intact."""
    from distributed_training_dl_b200.utils import PinnedPrefetcher
    g = torch.Generator().manual_seed(1)
    batches = [(torch.randn(64, 3, 32, 32, generator=g), torch.randint(0, 10, (64,), generator=g)) for _ in range(40)]
    dev_x = torch.empty(40, 64, 3, 32, 32, device="cuda")
    dev_t = torch.empty(40, 64, dtype=torch.long, device="cuda")
    ptrs = set()
    for i, (x, t) in enumerate(PinnedPrefetcher(batches, depth=3)):
        assert x.is_pinned() and t.is_pinned()
        ptrs.add(x.data_ptr())
        torch.cuda._sleep(2_000_000)                 # keep the stream busy so copies really lag the host
        dev_x[i].copy_(x, non_blocking=True)
        dev_t[i].copy_(t, non_blocking=True)
    torch.cuda.synchronize()
    assert len(ptrs) <= 3
    for i, (x, t) in enumerate(batches):
        assert torch.equal(dev_x[i].cpu(), x) and torch.equal(dev_t[i].cpu(), t), i


def test_weights_loaded_into_a_captured_step_are_used_by_the_next_replay():
    """A captured step re-packs the weight operands at its end, so weights loaded from outside between
    two replays used to be multiplied from the previous packs for one step.  TrainStep now notices the
    change (tensor version) and re-packs before replaying: a step after load_state_dict(zeros for the
    last layer) must produce exactly uniform logits, i.e. loss == log(10)."""
    import math
    import distributed_training_dl_b200 as dtdl
    from distributed_training_dl_b200.engine import CrossEntropyLoss, TrainStep
    from distributed_training_dl_b200.models import MnistCNN
    torch.manual_seed(0)
    net = dtdl.wrap(MnistCNN().cuda())
    opt = dtdl.FusedSGD(net.parameters(), lr=0.0, momentum=0.0)
    step = TrainStep(net, opt, CrossEntropyLoss(), graph=True, warmup=1)
    x = torch.randn(32, 1, 28, 28).pin_memory()
    t = torch.randint(0, 10, (32,)).pin_memory()
    for _ in range(4):
        before = step(x, t).item()
    assert step.launches_per_step, "graph not captured"
    assert abs(before - math.log(10)) > 1e-3
    sd = net.state_dict()
    last_w = [k for k in sd if k.endswith("weight")][-1]
    last_b = last_w[:-6] + "bias"
    sd[last_w] = torch.zeros_like(sd[last_w])
    sd[last_b] = torch.zeros_like(sd[last_b])
    net.load_state_dict(sd)
    after = step(x, t).item()          # a replay; lr = 0, so nothing else changes
    assert abs(after - math.log(10)) < 1e-4, (before, after)


def test_keeping_the_returned_loss_does_not_break_the_capture():
    """``Model.fit`` keeps each step's loss until the next call returns.  The loss of an eager warm-up
    step used to carry that step's autograd graph, whose gradient-accumulation nodes are bound to the
    eager stream; the capture of the next call then failed with "dependency created on uncaptured
    work".  The step now returns a detached loss, with an optimizer whose hyper-parameters change
    every step (Adam) as well."""
    import distributed_training_dl_b200 as dtdl
    from distributed_training_dl_b200.engine import CrossEntropyLoss, TrainStep
    from distributed_training_dl_b200.models import MnistCNN
    torch.manual_seed(0)
    net = dtdl.wrap(MnistCNN().cuda())
    opt = dtdl.FusedAdam(net.parameters(), lr=1e-3, keras=True)
    step = TrainStep(net, opt, CrossEntropyLoss(), graph=True, warmup=2)
    g = torch.Generator().manual_seed(1)
    x = torch.randn(32, 1, 28, 28, generator=g).pin_memory()
    t = torch.randint(0, 10, (32,), generator=g).pin_memory()
    kept, losses = None, []
    for _ in range(6):
        kept = step(x, t)               # the previous ``kept`` is alive during this call
        assert kept.grad_fn is None
        losses.append(kept.item())
    assert step.launches_per_step and step.launches_per_step > 0      # the graph was captured
    assert losses[-1] < losses[0], losses                             # same batch: the loss goes down
