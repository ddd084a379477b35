"""Development aid: kernel-time breakdown of one eager PyramidNet training step (shares only)."""
import sys
import torch
sys.path.insert(0, ".")
import distributed_training_dl_b200 as dtdl
from distributed_training_dl_b200.engine import CrossEntropyLoss, TrainStep
from torch.profiler import profile, ProfilerActivity

comm = dtdl.init(backend="b200")
net = dtdl.wrap(dtdl.models.pyramidnet().cuda())
opt = dtdl.FusedSGD(net.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
step = TrainStep(net, opt, CrossEntropyLoss(), graph=False)
x = torch.randn(64, 3, 32, 32, device="cuda"); t = torch.randint(0, 10, (64,), device="cuda")
for _ in range(3):
    step(x, t)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    step(x, t)
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=25, max_name_column_width=60))
