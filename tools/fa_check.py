import os, sys
sys.path.insert(0, "/root/repo")
import torch
from torch.profiler import ProfilerActivity, profile
from b200ddp.ops import functional as Fn
dev = torch.device("cuda")
qkv = torch.randn(8, 1024, 3 * 768, device=dev).to(torch.bfloat16)
for _ in range(3): Fn.flash_attention_forward(qkv, 12)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(5): Fn.flash_attention_forward(qkv, 12)
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=5, max_name_column_width=60))
