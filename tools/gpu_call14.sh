#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
run() { local name=$1 t=$2; shift 2; echo "=== $name"; timeout "$t" "$@" > "gpurun_out/$name.log" 2>&1; echo "rc=$?"; grep -v "OMP_NUM_THREADS\|^\*\*\*\*" "gpurun_out/$name.log" | tail -n "${TAILN:-15}"; }
TAILN=4 run t_bn 600 python -m pytest tests/test_gpu_kernels.py -q -m gpu -k "batchnorm or resnet" --timeout 300 -p no:cacheprovider
TAILN=1 run rn50_ours 600 python bench/bench_models.py --model resnet50 --batch-size 64 --steps 10 --warmup 3 --pool 2
TAILN=12 run train_simplenet_2gpu 600 python -m b200ddp.launch --nproc-per-node=2 --master-port=29601 -m b200ddp.train --epochs 3 --num-samples 20000 --checkpoint-dir /tmp/ck_simple
TAILN=6 run train_simplenet_resume 600 python -m b200ddp.launch --nproc-per-node=2 --master-port=29602 -m b200ddp.train --epochs 4 --num-samples 20000 --checkpoint-dir /tmp/ck_simple --resume /tmp/ck_simple/latest_model.pt
TAILN=8 run train_gpt2_2gpu 900 python -m b200ddp.launch --nproc-per-node=2 --master-port=29603 -m b200ddp.train --model gpt2 --seq-len 256 --epochs 1 --num-samples 128 --batch-size 8 --optimizer adamw --lr 3e-4 --checkpoint-dir /tmp/ck_gpt2
TAILN=8 run train_rn50_1gpu 900 python -m b200ddp.launch --nproc-per-node=1 --master-port=29604 -m b200ddp.train --model resnet50 --epochs 1 --num-samples 256 --batch-size 32 --optimizer sgd --lr 0.05 --checkpoint-dir /tmp/ck_rn50
python - <<'PY'
import torch, sys
sys.path.insert(0, '.')
import importlib.util
spec = importlib.util.spec_from_file_location("ref", "baseline/_ref/train.py"); ref = importlib.util.module_from_spec(spec); spec.loader.exec_module(ref)
m = ref.SimpleNet(); o = torch.optim.Adam(m.parameters())
print("reference loads GPU-written checkpoint, epoch", ref.load_checkpoint(m, o, "/tmp/ck_simple/latest_model.pt", torch.device("cpu")))
PY
