"""Multi-GPU tier: peer-memory collectives and the engine at world size >= 2."""
import os
import subprocess
import sys

import pytest
import torch

from conftest import free_port

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_peer_memory_collectives_and_engine():
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs on the box")
    world = 8 if n >= 8 else (4 if n >= 4 else 2)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(free_port()),
           os.path.join(ROOT, "tests", "multi_gpu_worker.py")]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=600)
    log = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(log):  # keep the full transcript where the GPU-box harness collects it
        with open(os.path.join(log, "multi_gpu_worker.log"), "w") as f:
            f.write(r.stdout + "\n---- stderr ----\n" + r.stderr)
    if r.returncode != 0:
        lines = [l for l in (r.stdout + r.stderr).splitlines()
                 if "Error" in l or "assert" in l or "matches" in l or "rank0" in l]
        pytest.fail("multi-GPU worker failed:\n" + "\n".join(lines[-25:]))
    assert "MULTI_GPU_OK" in r.stdout


def test_dead_peer_is_contained(tmp_path):
    """SURVEY §5 fault containment: a rank that dies before a fused collective must not wedge
    the others.  The survivor's kernel spins on the dead peer's flag, hits its bound
    (B200DDP_PEER_TIMEOUT_MS), traps; that rank exits non-zero and the launcher ends the job."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs on the box")
    script = tmp_path / "dead_peer.py"
    script.write_text("""
import os, sys, time, torch
from b200ddp import DataParallel, SimpleNet, setup_distributed
rank, world, local = setup_distributed("nccl")
dev = torch.device("cuda", local)
eng = DataParallel(SimpleNet().to(dev), backend="b200", compute_dtype=torch.bfloat16)
x = torch.randn(64, 784, device=dev); y = torch.randint(0, 10, (64,), device=dev)
eng.train_step(x, y); torch.cuda.synchronize()
print(f"STEP_OK {rank}", flush=True)
if rank == 1:
    os._exit(3)                       # dies between steps
t0 = time.time()
try:
    eng.train_step(x, y); torch.cuda.synchronize()
except Exception as e:
    print("SURVIVOR_ERROR after %.1fs: %s" % (time.time() - t0, type(e).__name__), flush=True)
    os._exit(9)
print("UNEXPECTED_SUCCESS", flush=True)
""")
    env = {**os.environ, "PYTHONPATH": ROOT, "B200DDP_PEER_TIMEOUT_MS": "2000"}
    import time
    t0 = time.time()
    r = subprocess.run([sys.executable, "-m", "b200ddp.launch", "--nproc-per-node=2",
                        f"--master-port={free_port()}", str(script)],
                       cwd=ROOT, capture_output=True, text=True, timeout=180, env=env)
    assert r.returncode != 0
    assert "STEP_OK 0" in r.stdout and "STEP_OK 1" in r.stdout
    assert "UNEXPECTED_SUCCESS" not in r.stdout
    assert time.time() - t0 < 120
