"""Development aid: gradient cosine (GPU bf16 kernels vs CPU fp32) along PyramidNet's depth."""
import copy, sys
import torch
import torch.nn.functional as F
sys.path.insert(0, ".")
from distributed_training_dl_b200.models import PyramidNet

torch.manual_seed(0)
for layers, batch in [(3, 8), (18, 8), (18, 32)]:
    cpu = PyramidNet(num_layers=layers, alpha=270.0 * layers / 18); gpu = copy.deepcopy(cpu).cuda()
    x = torch.randn(batch, 3, 32, 32); t = torch.randint(0, 10, (batch,))
    F.cross_entropy(cpu(x), t).backward(); F.cross_entropy(gpu(x.cuda()), t.cuda()).backward()
    rows = []
    for (n, a), (_, b) in zip(cpu.named_parameters(), gpu.named_parameters()):
        if a.dim() != 4: continue
        ga, gb = a.grad.flatten().double(), b.grad.cpu().flatten().double()
        rows.append((n, (torch.dot(ga, gb) / (ga.norm() * gb.norm())).item(), (gb.norm() / ga.norm()).item()))
    print(f"layers={layers} batch={batch}")
    step = max(1, len(rows) // 12)
    for n, c, r in rows[::step] + rows[-1:]:
        print(f"   {n:28s} cos {c:.4f} norm_ratio {r:.3f}")
