#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
N=${1:-2}
run() { local name=$1 t=$2; shift 2; echo "=== $name"; timeout "$t" "$@" > "gpurun_out/$name.log" 2>&1; echo "rc=$?"; tail -n "${TAILN:-15}" "gpurun_out/$name.log"; }
TAILN=30 run t_all_gpu 1200 python -m pytest tests/ -x -q -m gpu --timeout 600 -p no:cacheprovider
TAILN=30 run multi_direct 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 tests/multi_gpu_worker.py
TAILN=4 run bench_ref_n 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29534 bench.py --impl reference --gpus $N --steps 500 --warmup 50
TAILN=4 run bench_ours_n 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29535 bench.py --gpus $N --steps 2000 --warmup 200
