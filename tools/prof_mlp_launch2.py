"""Per-launch device time of consecutive 64-step launches of the fused step (events between launches)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
os.environ.setdefault("RANK", "0"); os.environ.setdefault("WORLD_SIZE", "1")
os.environ.setdefault("MASTER_ADDR", "127.0.0.1"); os.environ.setdefault("MASTER_PORT", "29579")
torch.cuda.set_device(0)
dist.init_process_group("nccl", device_id=torch.device("cuda", 0))
from b200ddp import DataParallel, SimpleNet
dev = torch.device("cuda", 0)
torch.manual_seed(0)
eng = DataParallel(SimpleNet().to(dev), backend="b200", compute_dtype=torch.bfloat16)
P = int(os.environ.get("POOL", "720"))
xs = torch.randn(P, 64, 784, device=dev); ys = torch.randint(0, 10, (P, 64), device=dev)
eng.train_steps(xs, ys, 0, 200)
torch.cuda.synchronize()
first = 200
import time
N = int(os.environ.get("LAUNCHES", "12"))
for mode in ("events between launches", "back to back", "events between launches"):
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(N + 1)]
    torch.cuda.synchronize()
    time.sleep(float(os.environ.get("SLEEP", "0")))
    ev[0].record()
    for k in range(N):
        eng.train_steps(xs, ys, first % P, 64); first += 64
        if mode.startswith("events") or k == N - 1:
            ev[k + 1].record()
    torch.cuda.synchronize()
    if mode.startswith("events"):
        print(mode, [round(ev[k].elapsed_time(ev[k + 1]) * 1e3 / 64, 2) for k in range(N)])
    else:
        print(mode, round(ev[0].elapsed_time(ev[N]) * 1e3 / (64 * N), 2), f"us/step over {N} launches")
print("loss", eng.last_loss())
dist.destroy_process_group()
