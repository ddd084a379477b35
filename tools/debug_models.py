"""Development aid: per-parameter gradient error of the GPU kernels vs the CPU fp32 oracle."""
import copy, sys
import torch
import torch.nn.functional as F
sys.path.insert(0, ".")
from distributed_training_dl_b200.models import MLP, MnistCNN, pyramidnet

torch.manual_seed(0)
for name, make, shape in [("mlp", lambda: MLP(128), (32, 784)), ("cnn", MnistCNN, (16, 1, 28, 28)), ("pyr", pyramidnet, (8, 3, 32, 32))]:
    cpu = make(); gpu = copy.deepcopy(cpu).cuda()
    x = torch.randn(*shape); t = torch.randint(0, 10, (shape[0],))
    oc = cpu(x); og = gpu(x.cuda())
    print(name, "logit err", (oc - og.cpu()).abs().max().item(), "ref max", oc.abs().max().item())
    lc = F.cross_entropy(oc, t); lg = F.cross_entropy(og, t.cuda())
    lc.backward(); lg.backward()
    errs = []
    for (n, a), (_, b) in zip(cpu.named_parameters(), gpu.named_parameters()):
        e = ((b.grad.cpu() - a.grad).abs().max() / (a.grad.abs().max() + 1e-12)).item()
        errs.append((e, n, a.grad.abs().max().item()))
    errs.sort(reverse=True)
    for e, n, m in errs[:6]:
        print(f"   {n:40s} rel_err {e:.4f} ref_max {m:.3e}")
