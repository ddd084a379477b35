"""Development aid: kernel-time breakdown of one eager PyramidNet training step (shares only).

    python tools/profile_step.py [bf16|tf32] [pyramidnet|resnet50]"""
import sys
import torch
sys.path.insert(0, ".")
import distributed_training_dl_b200 as dtdl
from distributed_training_dl_b200.engine import CrossEntropyLoss, TrainStep
from torch.profiler import profile, ProfilerActivity

comm = dtdl.init(backend="b200")
prec = sys.argv[1] if len(sys.argv) > 1 else "bf16"
model = sys.argv[2] if len(sys.argv) > 2 else "pyramidnet"
net = dtdl.wrap(getattr(dtdl.models, model)().cuda(), dtype=prec)
opt = dtdl.FusedSGD(net.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
step = TrainStep(net, opt, CrossEntropyLoss(), graph=False)
hw = 224 if model == "resnet50" else 32
x = torch.randn(64, 3, hw, hw, device="cuda"); t = torch.randint(0, 10, (64,), device="cuda")
print(f"{model} {prec}")
for _ in range(3):
    step(x, t)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    step(x, t)
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=22, max_name_column_width=70))
