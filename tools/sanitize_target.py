"""Small single-GPU workload for compute-sanitizer (memcheck / racecheck): every kernel family
once, tiny shapes."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b200ddp import DataParallel, SimpleNet
from b200ddp.models.gpt2 import GPT2, GPT2Config
from b200ddp.ops import functional as Fn
dev = torch.device("cuda")
torch.manual_seed(0)
eng = DataParallel(SimpleNet().to(dev), backend="b200", compute_dtype=torch.bfloat16)
x = torch.randn(64, 784, device=dev); y = torch.randint(0, 10, (64,), device=dev)
eng.program._train_body(eng.program._get(64)) if False else None
for _ in range(2):
    eng.train_step(x, y)
cfg = GPT2Config(vocab_size=512, n_positions=128, n_embd=256, n_layer=1, n_head=4)
g = DataParallel(GPT2(cfg).to(dev), optimizer="adamw", backend="b200", compute_dtype=torch.bfloat16)
idx = torch.randint(0, 512, (2, 128), device=dev)
g.train_step(idx, idx)
xc = torch.randn(2, 64, 16, 16, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last).requires_grad_(True)
w = (torch.randn(64, 64, 3, 3, device=dev) * 0.05).to(torch.bfloat16).requires_grad_(True)
Fn.conv2d(xc, w, 1, 1).float().sum().backward()
torch.cuda.synchronize()
print("SANITIZE_TARGET_OK")
