#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
run() { local name=$1 t=$2; shift 2; echo "=== $name"; timeout "$t" "$@" > "gpurun_out/$name.log" 2>&1; echo "rc=$?"; grep -v "OMP_NUM_THREADS\|^\*\*\*\*" "gpurun_out/$name.log" | tail -n "${TAILN:-15}"; }
TAILN=25 run t_conv 600 python -m pytest tests/test_gpu_kernels.py -q -m gpu -k "conv2d or resnet" --timeout 300 -p no:cacheprovider
TAILN=3 run rn50_ours 900 python bench/bench_models.py --model resnet50 --batch-size 64 --steps 10 --warmup 3 --pool 2
