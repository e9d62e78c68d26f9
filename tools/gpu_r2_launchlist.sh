#!/usr/bin/env bash
# launch list (device time per kernel, caches left warm) of the fused step; PDL on/off comparison
set -u
mkdir -p gpurun_out
python tools/prof_mlp.py 60 > gpurun_out/ll_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -s 60 -c 40 --csv --log-file gpurun_out/ll_launches.csv python tools/prof_mlp.py 60 > gpurun_out/ll_ncu.log 2>&1
echo "ncu rc=$?"
python - <<'PY'
import csv
rows=[r for r in csv.reader(open("gpurun_out/ll_launches.csv")) if len(r)>5]
hdr=rows[0]; ki=hdr.index("Kernel Name"); vi=hdr.index("Metric Value")
for r in rows[1:9]: print(r[ki][:70], r[vi])
PY
echo "--- PDL on"; head -4 gpurun_out/ll_plain.log | tail -3
echo "--- PDL off"; B200DDP_NO_PDL=1 python tools/prof_mlp.py 200 2>&1 | head -4 | tail -3
