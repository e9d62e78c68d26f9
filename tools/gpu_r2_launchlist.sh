#!/usr/bin/env bash
# launch list (device time per kernel) of the fused step, then the whole GPU tier
set -u
mkdir -p gpurun_out
python tools/prof_mlp.py 60 > gpurun_out/ll_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 60 -c 40 --csv --log-file gpurun_out/ll_launches.csv python tools/prof_mlp.py 60 > gpurun_out/ll_ncu.log 2>&1
echo "ncu rc=$?"
python - <<'PY'
import csv
rows=[r for r in csv.reader(open("gpurun_out/ll_launches.csv")) if len(r)>5]
hdr=rows[0]; ki=hdr.index("Kernel Name"); vi=hdr.index("Metric Value")
for r in rows[1:21]: print(r[ki][:70], r[vi])
PY
timeout 900 python -m pytest tests -q -m gpu -x --timeout 300 -p no:cacheprovider > gpurun_out/ll_tests.log 2>&1; echo "tests rc=$?"; tail -5 gpurun_out/ll_tests.log
