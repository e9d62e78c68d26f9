#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
run() { local name=$1 t=$2; shift 2; echo "=== $name"; timeout "$t" "$@" > "gpurun_out/$name.log" 2>&1; echo "rc=$?"; grep -v "OMP_NUM_THREADS\|^\*\*\*\*" "gpurun_out/$name.log" | tail -n "${TAILN:-15}"; }
TAILN=25 run multi_nvls_n$N 600 $TR --master-port 29533 tests/multi_gpu_worker.py

TAILN=14 run sweep_bf16_n$N 600 $TR --master-port 29541 bench/allreduce_sweep.py --dtype bf16 --out gpurun_out/sweep_bf16_n$N.json
TAILN=2 run bench_ours_n$N 300 $TR --master-port 29535 bench.py --gpus $N --steps 2000 --warmup 200
B200DDP_ONE_SHOT_MAX_BYTES=0 TAILN=2 run bench_ours_n${N}_nvls_forced 300 $TR --master-port 29536 bench.py --gpus $N --steps 2000 --warmup 200 --no-e2e
