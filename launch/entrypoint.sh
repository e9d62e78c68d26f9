#!/usr/bin/env bash
# Container entrypoint: derive this pod's node rank and the rendezvous address
# from a StatefulSet-style hostname, then hand over to the rank supervisor.
#
# Environment contract (identical to the reference, entrypoint.sh:5-8,19,39):
#   NF_DISCOVERY_SERVICE  headless-service DNS suffix        (required when REPLICAS > 1)
#   REPLICAS              number of nodes                    (required)
#   NPROC_PER_NODE        workers per node, or "gpu"         (default 1)
#   MASTER_PORT           rendezvous port                    (default 29500)
#   TRAINING_SCRIPT       script run by every worker         (default: the b200ddp trainer)
#   SCRIPT_ARGS           extra trainer arguments, word-split on purpose
# Additions:
#   LAUNCHER=torchrun     use torchrun instead of b200ddp.launch
#   single-box mode: REPLICAS=1 needs no discovery service; master is 127.0.0.1
#   MAX_RESTARTS=N        (b200ddp.launch, single node) restart the worker group up to N times;
#   RESTART_ARGS="..."    trainer arguments added on restarts, e.g. "--resume ck/latest_model.pt"
set -euo pipefail

say() { printf '[%s] %s\n' "$(date -u +%H:%M:%S)" "$*"; }
die() { say "ERROR: $*"; exit 1; }

port="${MASTER_PORT:-29500}"
nproc="${NPROC_PER_NODE:-1}"
script="${TRAINING_SCRIPT:-}"
launcher="${LAUNCHER:-b200ddp}"

[ -n "${REPLICAS:-}" ] || die "REPLICAS not set"

if [ "${REPLICAS}" = "1" ] && [ -z "${NF_DISCOVERY_SERVICE:-}" ]; then
    node_rank=0
    master="127.0.0.1"
else
    [ -n "${NF_DISCOVERY_SERVICE:-}" ] || die "NF_DISCOVERY_SERVICE not set"
    host="$(hostname)"
    node_rank="${host##*-}"
    case "${node_rank}" in
        ''|*[!0-9]*) die "hostname '${host}' does not end in -<ordinal>" ;;
    esac
    master="${host%-*}-0.${NF_DISCOVERY_SERVICE}"
fi

say "Node ${node_rank}/${REPLICAS} starting, master=${master}"

topology=(--nnodes="${REPLICAS}" --nproc-per-node="${nproc}" --node-rank="${node_rank}"
          --master-addr="${master}" --master-port="${port}")

if [ "${launcher}" = "torchrun" ]; then
    say "Starting torchrun (rendezvous will synchronize nodes)"
    # shellcheck disable=SC2086
    exec torchrun "${topology[@]}" ${script:-train.py} ${SCRIPT_ARGS:-}
fi

say "Starting b200ddp.launch (static rendezvous, ${nproc} worker(s) on this node)"
if [ -n "${MAX_RESTARTS:-}" ]; then
    topology+=(--max-restarts="${MAX_RESTARTS}")
    [ -z "${RESTART_ARGS:-}" ] || topology+=(--restart-args="${RESTART_ARGS}")
fi
if [ -n "${script}" ]; then
    # shellcheck disable=SC2086
    exec python -m b200ddp.launch "${topology[@]}" "${script}" ${SCRIPT_ARGS:-}
fi
# shellcheck disable=SC2086
exec python -m b200ddp.launch "${topology[@]}" -m b200ddp.train ${SCRIPT_ARGS:-}
