#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
run() { local name=$1 t=$2; shift 2; echo "=== $name"; timeout "$t" "$@" > "gpurun_out/$name.log" 2>&1; echo "rc=$?"; grep -v "OMP_NUM_THREADS\|^\*\*\*\*" "gpurun_out/$name.log" | tail -n "${TAILN:-15}"; }
TAILN=1 run n2b_bench 300 $TR --nproc-per-node 2 --master-port 29716 bench.py --gpus 2 --steps 1000 --warmup 100
TAILN=6 run n2b_train 300 $TR --nproc-per-node 2 --master-port 29717 train.py --epochs 2 --num-samples 20000 --checkpoint-dir /tmp/ck
TAILN=5 run n2b_train_rn 300 $TR --nproc-per-node 2 --master-port 29718 train.py --epochs 1 --num-samples 256 --model resnet50 --batch-size 16 --checkpoint-dir /tmp/ck3
