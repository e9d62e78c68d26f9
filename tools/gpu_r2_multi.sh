#!/usr/bin/env bash
# Multi-GPU pass: numerics worker, bench at N GPUs (ours + reference), all-reduce sweep.
# usage: bash tools/gpu_r2_multi.sh N [sweep]
set -u
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node $N"
run() { local name=$1 t=$2; shift 2; echo "=== $name"; timeout "$t" "$@" > "gpurun_out/$name.log" 2>&1; echo "rc=$?"; grep -v "OMP_NUM_THREADS\|^\*\*\*\*\|^$" "gpurun_out/$name.log" | tail -n "${TAILN:-12}"; }
TAILN=30 run m${N}_worker 500 $TR --master-port 29701 tests/multi_gpu_worker.py
TAILN=1 run m${N}_bench20 200 $TR --master-port 29702 bench.py --gpus $N --steps 20 --warmup 5
TAILN=1 run m${N}_bench2000 200 $TR --master-port 29703 bench.py --gpus $N --steps 2000 --warmup 200
if [ "${2:-}" = "sweep" ]; then
  TAILN=40 run m${N}_sweep_bf16 400 $TR --master-port 29704 bench/allreduce_sweep.py --dtype bf16
fi
