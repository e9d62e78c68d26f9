#!/usr/bin/env bash
# quick 1-GPU model benches (GPT-2 small, ResNet-50, train.py model)
set -u
mkdir -p gpurun_out
timeout 300 python bench/bench_models.py --model gpt2 --steps 10 --warmup 3 2>&1 | grep '^{' | tee gpurun_out/q_gpt2.json | cut -c1-330
timeout 300 python bench/bench_models.py --model resnet50 --batch-size 64 --steps 10 --warmup 3 2>&1 | grep '^{' | tee gpurun_out/q_rn50.json | cut -c1-330
timeout 300 python bench.py --steps 200 --warmup 20 2>&1 | grep '^{' | tee gpurun_out/q_bench.json | cut -c1-400
