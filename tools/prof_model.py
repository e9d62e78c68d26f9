"""Kernel-time breakdown of one training step of a model config (torch.profiler, CUDA activities)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import ProfilerActivity, profile
from b200ddp import DataParallel
model = sys.argv[1] if len(sys.argv) > 1 else "resnet50"
bsz = int(sys.argv[2]) if len(sys.argv) > 2 else 64
dev = torch.device("cuda")
torch.manual_seed(0)
if model == "resnet50":
    from b200ddp.models.resnet50 import resnet50
    net = resnet50().to(dev)
    x = torch.randn(bsz, 3, 224, 224, device=dev); y = torch.randint(0, 1000, (bsz,), device=dev)
else:
    from b200ddp.models.gpt2 import GPT2
    net = GPT2().to(dev)
    t = torch.randint(0, 50257, (bsz, 1025), device=dev); x, y = t[:, :-1].contiguous(), t[:, 1:].contiguous()
eng = DataParallel(net, optimizer="adamw", lr=3e-4, backend="b200", compute_dtype=torch.bfloat16)
for _ in range(3):
    eng.train_step(x, y)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(2):
        eng.train_step(x, y)
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=28, max_name_column_width=70))
