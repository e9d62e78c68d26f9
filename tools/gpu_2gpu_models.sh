#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
run() { local name=$1 t=$2; shift 2; echo "=== $name"; timeout "$t" "$@" > "gpurun_out/$name.log" 2>&1; echo "rc=$?"; grep -v "OMP_NUM_THREADS\|^\*\*\*\*" "gpurun_out/$name.log" | tail -n "${TAILN:-15}"; }
TAILN=14 run n2_multi_tests 600 python -m pytest tests/test_gpu_multi.py -q -m gpu --timeout 500 -p no:cacheprovider
TAILN=1 run n2_gpt2 300 $TR --nproc-per-node 2 --master-port 29714 bench/bench_models.py --model gpt2 --gpus 2 --batch-size 8 --steps 10 --warmup 3
TAILN=1 run n2_rn50 300 $TR --nproc-per-node 2 --master-port 29715 bench/bench_models.py --model resnet50 --gpus 2 --batch-size 64 --steps 10 --warmup 3 --pool 2
TAILN=1 run n2_bench 300 $TR --nproc-per-node 2 --master-port 29716 bench.py --gpus 2 --steps 1000 --warmup 100
