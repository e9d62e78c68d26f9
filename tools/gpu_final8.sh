#!/usr/bin/env bash
# Final multi-GPU validation (8 GPUs): numerics, containment, scaling points, model configs.
set -u
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
run() { local name=$1 t=$2; shift 2; echo "=== $name"; timeout "$t" "$@" > "gpurun_out/$name.log" 2>&1; echo "rc=$?"; grep -v "OMP_NUM_THREADS\|^\*\*\*\*" "gpurun_out/$name.log" | tail -n "${TAILN:-15}"; }
TAILN=22 run f_multi_tests 900 python -m pytest tests/test_gpu_multi.py -q -m gpu --timeout 800 -p no:cacheprovider
TAILN=1 run f_bench_ours_n4 300 $TR --nproc-per-node 4 --master-port 29701 bench.py --gpus 4 --steps 2000 --warmup 200
TAILN=1 run f_bench_ref_n4 400 $TR --nproc-per-node 4 --master-port 29702 bench.py --impl reference --gpus 4 --steps 300 --warmup 30
TAILN=1 run f_bench_ours_n8 300 $TR --nproc-per-node 8 --master-port 29703 bench.py --gpus 8 --steps 2000 --warmup 200
TAILN=1 run f_gpt2_ours_n8 600 $TR --nproc-per-node 8 --master-port 29704 bench/bench_models.py --model gpt2 --gpus 8 --batch-size 8 --steps 10 --warmup 3
TAILN=1 run f_rn50_ours_n8 600 $TR --nproc-per-node 8 --master-port 29705 bench/bench_models.py --model resnet50 --gpus 8 --batch-size 64 --steps 10 --warmup 3 --pool 2
TAILN=1 run f_rn50_ref_n8 600 $TR --nproc-per-node 8 --master-port 29706 bench/bench_models.py --model resnet50 --gpus 8 --batch-size 64 --steps 10 --warmup 3 --pool 2 --impl reference
