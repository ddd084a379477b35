"""Tiny run of every single-GPU kernel for `compute-sanitizer --tool memcheck` (one tool per call)."""
import sys
import torch
sys.path.insert(0, ".")
import distributed_training_dl_b200 as dtdl
from distributed_training_dl_b200.models import PyramidNet, MnistCNN, MLP

torch.manual_seed(0)
comm = dtdl.init(backend="b200")
for make, shape in [(lambda: PyramidNet(num_layers=2, alpha=30.0), (4, 3, 32, 32)), (MnistCNN, (4, 1, 28, 28)), (lambda: MLP(64), (8, 784))]:
    net = dtdl.wrap(make().cuda())
    opt = dtdl.FusedSGD(net.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-4)
    crit = dtdl.CrossEntropyLoss()
    x = torch.randn(*shape, device="cuda"); t = torch.randint(0, 10, (shape[0],), device="cuda")
    for _ in range(2):
        loss = crit(net(x), t); opt.zero_grad(); loss.backward(); opt.step()
    torch.cuda.synchronize()
    print("ok", float(loss))
dtdl.shutdown()
