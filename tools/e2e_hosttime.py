"""Host time spent inside HostStepper.step (excluding the wait for the previous loss is not
possible from outside, so this reports: call time with the GPU idle vs. in steady state)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b200ddp import DataParallel, SimpleNet
dev = torch.device("cuda")
eng = DataParallel(SimpleNet().to(dev), optimizer="adam", lr=1e-3, backend="b200", compute_dtype=torch.bfloat16)
n = 512
hx = torch.randn(n, 64, 784).pin_memory(); hy = torch.randint(0, 10, (n, 64)).pin_memory()
st = eng.host_stepper()
for i in range(200):
    st.step(hx[i % n], hy[i % n])
st.finish(); torch.cuda.synchronize()
# steady state
t0 = time.perf_counter()
for i in range(4000):
    st.step(hx[i % n], hy[i % n])
st.finish(); torch.cuda.synchronize()
t1 = time.perf_counter()
print(f"steady state: {(t1 - t0) / 4000 * 1e6:.1f} us/step wall")
# GPU idle each time: pure host cost of one call (the previous loss is already there)
tot = 0.0
for i in range(500):
    torch.cuda.synchronize()
    a = time.perf_counter()
    st.step(hx[i % n], hy[i % n])
    tot += time.perf_counter() - a
print(f"host cost of step() with an idle GPU: {tot / 500 * 1e6:.1f} us")
xs = hx[:64].to(dev); ys = hy[:64].to(dev)
torch.cuda.synchronize()
t0 = time.perf_counter()
for i in range(4000):
    eng.train_step(xs[i % 64], ys[i % 64])
torch.cuda.synchronize()
print(f"train_step on device tensors: {(time.perf_counter() - t0) / 4000 * 1e6:.1f} us/step wall")
tot = 0.0
for i in range(500):
    torch.cuda.synchronize()
    a = time.perf_counter()
    hx[i % n]; hy[i % n]
    tot += time.perf_counter() - a
print(f"two pinned-tensor index ops: {tot / 500 * 1e6:.1f} us")
