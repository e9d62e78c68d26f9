#!/usr/bin/env bash
# 8-GPU (or N-GPU) confirmation: numerics worker with and without NVLS, bench (ours + reference)
set -u
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node $N"
run() { local name=$1 t=$2; shift 2; echo "=== $name"; timeout "$t" "$@" > "gpurun_out/$name.log" 2>&1; echo "rc=$?"; grep -v "OMP_NUM_THREADS\|^\*\*\*\*\|^$" "gpurun_out/$name.log" | tail -n "${TAILN:-12}"; }
TAILN=30 run m${N}_worker 500 $TR --master-port 29701 tests/multi_gpu_worker.py
TAILN=30 B200DDP_NVLS=0 run m${N}_worker_p2p 500 $TR --master-port 29705 tests/multi_gpu_worker.py
TAILN=1 run m${N}_bench20 200 $TR --master-port 29702 bench.py --gpus $N --steps 20 --warmup 5
TAILN=1 run m${N}_bench2000 200 $TR --master-port 29703 bench.py --gpus $N --steps 2000 --warmup 200
TAILN=1 B200DDP_ONE_SHOT_MAX_BYTES=100000000 run m${N}_bench2000_oneshot 200 $TR --master-port 29706 bench.py --gpus $N --steps 2000 --warmup 200 --no-e2e
TAILN=1 run m${N}_ref20 300 $TR --master-port 29704 bench.py --impl reference --gpus $N --steps 20 --warmup 5
