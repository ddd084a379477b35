import sys, traceback, torch
sys.path.insert(0, ".")
import distributed_training_dl_b200 as dtdl
from distributed_training_dl_b200.models import PyramidNet, MLP
from distributed_training_dl_b200.parallel import data_parallel as dpm

mode = sys.argv[1] if len(sys.argv) > 1 else "a"
ndev = torch.cuda.device_count()
torch.manual_seed(0)
dp = dtdl.DataParallel(MLP(64) if "mlp" in mode else PyramidNet(num_layers=3, alpha=45.0))
opt = dtdl.FusedSGD(dp.parameters(), lr=0.05, momentum=0.9)
import torch.nn.functional as F
if "pre" in mode or "ev" in mode:
    xd = torch.randn(8 * ndev, 3, 32, 32, device="cuda:0"); td = torch.randint(0, 10, (8 * ndev,), device="cuda:0")
    if "ev" in mode:
        dp.eval()
        with torch.no_grad(): dp(xd)
        dp.train()
    if "pre" in mode:
        for _ in range(2):
            loss = F.cross_entropy(dp(xd), td); opt.zero_grad(); loss.backward(); opt.step()
        if "del" in mode: del loss
    for d in range(ndev): torch.cuda.synchronize(d)
    print("pre ok")
step = dtdl.DataParallelStep(dp, opt, graph=False)
shape = (8 * ndev, 784) if "mlp" in mode else (8 * ndev, 3, 32, 32)
x = torch.randn(*shape).pin_memory(); t = torch.randint(0, 10, (8 * ndev,)).pin_memory()
for _ in range(2):
    step(x, t)
for d in range(ndev): torch.cuda.synchronize(d)
print("eager ok")
for g, dev in enumerate(dp.devices):
    try:
        with torch.cuda.device(dev):
            gr = torch.cuda.CUDAGraph()
            kw = {}
            if "s" in mode: kw["stream"] = step._streams[g]
            if "t" in mode: kw["capture_error_mode"] = "thread_local"
            with torch.cuda.graph(gr, **kw):
                if "nobwd" in mode:
                    m = dp
                    lo, hi = m.shard_bounds(step._sx.shape[0])[g]
                    out = m.run_replica_body(g, step._sx[lo:hi])
                else:
                    step._replica_step(g)
        print("capture ok on", dev)
    except Exception as e:
        print("capture FAILED on", dev, str(e).split("\n")[0])
        break
