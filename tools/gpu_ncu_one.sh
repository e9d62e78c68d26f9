#!/usr/bin/env bash
# usage: tools/gpu_ncu_one.sh <prof_target> <kernel regex> <out name>
set -u
mkdir -p gpurun_out
timeout 200 python tools/prof_targets.py $1 > gpurun_out/plain_$1.log 2>&1 || { echo "plain $1 failed"; tail -3 gpurun_out/plain_$1.log; exit 1; }
timeout 600 ncu --set full --clock-control none --import-source on -k regex:$2 -s 2 -c 1 -f -o gpurun_out/$3 python tools/prof_targets.py $1 > gpurun_out/ncu_$1.log 2>&1; echo rc=$?
tail -3 gpurun_out/ncu_$1.log
