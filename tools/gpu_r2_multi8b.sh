#!/usr/bin/env bash
# N-GPU confirmation of the single-launch step (update CTAs inside the step kernel): numerics worker,
# bench at 20 / 2000 steps, and the same bench with the update as a separate launch / forced one-shot.
set -u
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node $N"
run() { local name=$1 t=$2; shift 2; echo "=== $name"; timeout "$t" "$@" > "gpurun_out/$name.log" 2>&1; echo "rc=$?"; grep -v "OMP_NUM_THREADS\|^\*\*\*\*\|^$" "gpurun_out/$name.log" | tail -n "${TAILN:-12}" | cut -c1-${CUT:-400}; }
TAILN=12 run m${N}b_worker 400 $TR --master-port 29701 tests/multi_gpu_worker.py
CUT=1700 TAILN=1 run m${N}b_bench2000 200 $TR --master-port 29703 bench.py --gpus $N --steps 2000 --warmup 200
CUT=1700 TAILN=1 run m${N}b_bench20 200 $TR --master-port 29702 bench.py --gpus $N --steps 20 --warmup 5
TAILN=1 CUT=260 B200DDP_MLP_UPDATE_CLUSTERS=0 run m${N}b_bench2000_separate 200 $TR --master-port 29707 bench.py --gpus $N --steps 2000 --warmup 200 --no-e2e
TAILN=1 CUT=260 B200DDP_ONE_SHOT_MAX_BYTES=100000000 run m${N}b_bench2000_oneshot 200 $TR --master-port 29706 bench.py --gpus $N --steps 2000 --warmup 200 --no-e2e
