#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
for t in gemm256 flash step; do
  timeout 200 python tools/prof_targets.py $t > gpurun_out/plain_$t.log 2>&1 || { echo "plain $t failed"; tail -3 gpurun_out/plain_$t.log; exit 1; }
done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_kernel_m256 -s 2 -c 1 -o gpurun_out/prof_gemm_m256 python tools/prof_targets.py gemm256 > gpurun_out/ncu_gemm256.log 2>&1; echo rc=$?
timeout 600 ncu --set full --clock-control none --import-source on -k regex:flash_attn_fwd -s 2 -c 1 -o gpurun_out/prof_flash python tools/prof_targets.py flash > gpurun_out/ncu_flash.log 2>&1; echo rc=$?
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_step.csv python tools/prof_targets.py step > gpurun_out/ncu_step.log 2>&1; echo rc=$?
ls -la gpurun_out | tail -8
