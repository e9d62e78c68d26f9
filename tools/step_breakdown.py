"""Where does the SimpleNet step time go: device graph vs host-side Python/launch overhead."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b200ddp import DataParallel, SimpleNet
from b200ddp._native import bindings as B
dev = torch.device("cuda")
eng = DataParallel(SimpleNet().to(dev), backend="b200", compute_dtype=torch.bfloat16)
x = torch.randn(64, 784, device=dev); y = torch.randint(0, 10, (64,), device=dev)
for _ in range(50): eng.train_step(x, y)
bufs = eng.program._get(64)
def timeit(fn, n=2000):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3, (time.perf_counter() - t0) / n * 1e6
print("graph replay only        : %.1f us device, %.1f us wall" % timeit(lambda: bufs.train_graph.replay()))
print("stage + replay           : %.1f us device, %.1f us wall" % timeit(lambda: (B.stage_batch(x, y, bufs.x, bufs.y, bufs.zero_train), bufs.train_graph.replay())))
print("program.train_step       : %.1f us device, %.1f us wall" % timeit(lambda: eng.program.train_step(x, y)))
print("engine.train_step        : %.1f us device, %.1f us wall" % timeit(lambda: eng.train_step(x, y)))
t0 = time.perf_counter()
for _ in range(2000): bufs.train_graph.replay()
t1 = time.perf_counter(); torch.cuda.synchronize()
print("host cost of replay() alone (no sync): %.1f us" % ((t1 - t0) / 2000 * 1e6))
