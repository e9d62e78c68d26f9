"""Fixed cost of one fused-step launch: device time of train_steps(n) for several n, synchronised
before each call (what bench.py --steps 20 sees).   python tools/prof_mlp_launch.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
os.environ.setdefault("RANK", "0"); os.environ.setdefault("WORLD_SIZE", "1")
os.environ.setdefault("MASTER_ADDR", "127.0.0.1"); os.environ.setdefault("MASTER_PORT", "29578")
LOCAL = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(LOCAL)
dist.init_process_group("nccl", device_id=torch.device("cuda", LOCAL))
from b200ddp import DataParallel, SimpleNet
dev = torch.device("cuda", LOCAL)
torch.manual_seed(0)
eng = DataParallel(SimpleNet().to(dev), backend="b200", compute_dtype=torch.bfloat16)
xs = torch.randn(720, 64, 784, device=dev); ys = torch.randint(0, 10, (720, 64), device=dev)
eng.train_steps(xs, ys, 0, 200)
torch.cuda.synchronize()
first = 200
for n in [int(v) for v in os.environ.get("NS", "1,2,5,10,20,40,64,65,128,640").split(",")]:
    best = 1e9
    for rep in range(5):
        dist.barrier(); torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); eng.train_steps(xs, ys, first % 720, n); b.record(); torch.cuda.synchronize()
        first += n
        best = min(best, a.elapsed_time(b) * 1e3)
    if dist.get_rank() == 0:
        print(f"n = {n:4d}: {best:9.1f} us total, {best / n:7.2f} us/step")
dist.destroy_process_group()
