#!/usr/bin/env bash
# Late-round 8-GPU confirmation: multi-GPU numerics (NVLS inside graph replay), headline bench, model configs.
set -u
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
run() { local name=$1 t=$2; shift 2; echo "=== $name"; timeout "$t" "$@" > "gpurun_out/$name.log" 2>&1; echo "rc=$?"; grep -v "OMP_NUM_THREADS\|^\*\*\*\*" "gpurun_out/$name.log" | tail -n "${TAILN:-15}"; }
TAILN=4 run g_multi_tests 400 python -m pytest tests/test_gpu_multi.py -q -m gpu --timeout 350 -p no:cacheprovider
TAILN=1 run g_bench_ours_n8 200 $TR --nproc-per-node 8 --master-port 29703 bench.py --gpus 8 --steps 2000 --warmup 200
TAILN=1 run g_gpt2_ours_n8 300 $TR --nproc-per-node 8 --master-port 29704 bench/bench_models.py --model gpt2 --gpus 8 --batch-size 8 --steps 10 --warmup 4
TAILN=1 run g_rn50_ours_n8 300 $TR --nproc-per-node 8 --master-port 29705 bench/bench_models.py --model resnet50 --gpus 8 --batch-size 64 --steps 10 --warmup 4 --pool 2
