"""Timing / ncu target for the fused SimpleNet step: per-kernel device time, start-up transient and
(B200DDP_MLP_TRACE=1) the in-kernel phase timeline.
    python tools/prof_mlp.py [steps]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
os.environ.setdefault("RANK", "0"); os.environ.setdefault("WORLD_SIZE", "1")
os.environ.setdefault("MASTER_ADDR", "127.0.0.1"); os.environ.setdefault("MASTER_PORT", "29577")
LOCAL = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(LOCAL)
dist.init_process_group("nccl", device_id=torch.device("cuda", LOCAL))
if dist.get_rank() != 0:
    sys.stdout = open(os.devnull, "w")
from b200ddp import DataParallel, SimpleNet
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 200
dev = torch.device("cuda", LOCAL)
torch.manual_seed(0)
eng = DataParallel(SimpleNet().to(dev), backend="b200", compute_dtype=torch.bfloat16)
xs = torch.randn(64, 64, 784, device=dev); ys = torch.randint(0, 10, (64, 64), device=dev)
fs = eng.program.fused
def ev(): return torch.cuda.Event(enable_timing=True)
def timeit(fn, n):
    torch.cuda.synchronize(); a, b = ev(), ev(); a.record(); fn(n); b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) * 1e3 / n
eng.train_steps(xs, ys, 0, 20)
print("full step              us:", round(timeit(lambda n: eng.train_steps(xs, ys, 0, n), steps), 2))
def fb(train):
    def f(n):
        for i in range(n): fs.forward_backward(xs[i % 64], ys[i % 64], train=train)
    return f
print("fwd+bwd kernel only    us:", round(timeit(fb(True), steps), 2))
print("fwd-only (eval) kernel us:", round(timeit(fb(False), steps), 2))
import time; time.sleep(1.0)
marks = [ev() for _ in range(9)]
marks[0].record()
for k in range(8):
    eng.train_steps(xs, ys, 0, 5); marks[k + 1].record()
torch.cuda.synchronize()
print("us/step in chunks of 5 after 1 s idle:", [round(marks[k].elapsed_time(marks[k + 1]) * 200, 1) for k in range(8)])
if os.environ.get("B200DDP_MLP_TRACE"):
    import ctypes as C
    from b200ddp import _native as nat
    if fs.update_clusters:
        print("update clusters in the step launch:", fs.update_clusters)
        eng.train_steps(xs, ys, 0, 40)
        ub = (C.c_ulonglong * 34)()
        nat.lib().b200_mlp_debug_utrace.argtypes = [C.c_void_p]
        nat.check(nat.lib().b200_mlp_debug_utrace(ub), "utrace")
        un = ["compute entry", "compute past wait", "compute signal", "upd0 grads loaded", "upd0 entry", "upd0 past wait",
              "upd0 scalars", "upd0 flag seen", "upd0 loop done", "upd0 stores issued", "updL entry", "updL past wait",
              "updL scalars", "updL flag seen", "updL loop done", "-"]
        a, b = (ub[0:16], ub[16:32]) if ub[0] < ub[16] else (ub[16:32], ub[0:16])
        t0 = a[0]
        print("merged-update wall clock (globaltimer ns since the earlier launch's entry): earlier launch | later launch")
        for i, nm in enumerate(un):
            if nm != "-" and a[i] and abs(a[i] - t0) < 10**9:     # marks this path does not write stay zero
                print(f"  {nm:18s} {a[i] - t0:8d} {b[i] - t0:8d}")
        print(f"  last update CTA counted in (any step) {ub[32] - t0:8d}; compute saw the update (last time) {ub[33] - t0:8d}")
    names = ["entry", "x staged", "cta sync", "P1 mma done", "P1 drained", "cluster sync 1", "F1 done", "P2 mma done",
             "P2 drained", "cluster sync 2", "F2 done", "P3 done", "cluster sync 3", "CE done", "P4 mma done",
             "P4 epilogue", "P5 mma done", "P5 drained", "cluster sync 4", "F5 done", "P6 mma done", "P6 stored", "exit",
             "P4 gate done"]
    order = list(range(15)) + [23] + list(range(15, 23))
    for train in (True, False):
        fs.forward_backward(xs[3], ys[3], train=train)
        buf = (C.c_longlong * (8 * 24))()
        nat.lib().b200_mlp_debug_trace.argtypes = [C.c_void_p]
        nat.check(nat.lib().b200_mlp_debug_trace(buf), "trace")
        print("phase timeline, train =", train, "(us since entry at 1.965 GHz; CTA 0 .. 7)")
        for i in order:
            if not train and 13 < i < 21 or (not train and i == 23):
                continue
            print(f"  {names[i]:15s}", " ".join(f"{(buf[c * 24 + i] - buf[c * 24]) / 1965.0:7.2f}" for c in range(8)))
dist.destroy_process_group()
