"""Is the low fp32-vs-bf16 gradient cosine of random-init ResNet-50 inherent?  Stock torchvision
ResNet-50 on the GPU: fp32 vs bf16 autocast, same weights and inputs."""
import copy, torch, torchvision
import torch.nn.functional as F
torch.manual_seed(0)
a = torchvision.models.resnet50(num_classes=100).cuda(); b = copy.deepcopy(a)
x = torch.randn(8, 3, 128, 128, device="cuda"); t = torch.randint(0, 100, (8,), device="cuda")
torch.backends.cudnn.allow_tf32 = False; torch.backends.cuda.matmul.allow_tf32 = False
F.cross_entropy(a(x), t).backward()
with torch.autocast("cuda", dtype=torch.bfloat16):
    lb = F.cross_entropy(b(x), t)
lb.backward()
rows = []
for (n, p), (_, q) in zip(a.named_parameters(), b.named_parameters()):
    ga, gb = p.grad.flatten().double(), q.grad.flatten().double()
    rows.append((n, (torch.dot(ga, gb) / (ga.norm() * gb.norm() + 1e-30)).item()))
for n, c in rows[::16] + rows[-2:]:
    print(f"{n:36s} cos {c:.4f}")
