#!/usr/bin/env python3
"""Node launcher and rank supervisor (``python -m b200ddp.launch``).

Takes the place of ``torchrun`` in the reference's start path (reference
``entrypoint.sh:33-39``): static rendezvous on ``--master-addr:--master-port``,
``--nnodes × --nproc-per-node`` workers, ``--node-rank`` from the pod hostname.

Behaviour kept from torchrun as the reference uses it:

* every worker gets the elastic-agent environment contract — ``RANK``,
  ``LOCAL_RANK``, ``WORLD_SIZE``, ``LOCAL_WORLD_SIZE``, ``GROUP_RANK``,
  ``ROLE_RANK``, ``MASTER_ADDR``, ``MASTER_PORT``, ``TORCHELASTIC_*``,
  ``TORCH_NCCL_ASYNC_ERROR_HANDLING=1`` and ``OMP_NUM_THREADS=1`` when more than
  one worker shares the node and the variable is unset;
* workers are polled every 0.1 s; with the default ``--max-restarts 0`` (what the
  reference runs with) the first worker that exits non-zero fails the node, its
  siblings are terminated (SIGTERM, then SIGKILL after a grace period) and the
  launcher exits with that worker's code;
* ``--max-restarts N`` restarts the whole worker group up to N times, like torchrun's
  agent: ``TORCHELASTIC_RESTART_COUNT`` counts the attempt, the rendezvous port moves
  by one per attempt (the old TCPStore may linger in TIME_WAIT), and
  ``--restart-args`` appends arguments to the training script on restarts — e.g.
  ``--restart-args="--resume ck/latest_model.pt --resume-next-epoch"``.  Single node
  only: a static rendezvous cannot re-admit nodes that restart at different times.

What is different: there is no agent-side store or elastic machinery.  Global
rank 0 hosts the env:// TCPStore itself, and the supervisor is ~100 lines with
no dependencies, so a wedged rank cannot hide behind a multi-layer agent: peers
spinning on a dead rank's flags are torn down by the launcher (in-kernel spins
additionally carry their own timeout, see ``csrc/comm``).
"""
from __future__ import annotations

import argparse
import os
import shlex
import signal
import socket
import subprocess
import sys
import time
import uuid
from typing import Dict, List, Optional

MONITOR_INTERVAL_S = 0.1
TERM_GRACE_S = 5.0


def log(msg: str) -> None:
    print(f"[{time.strftime('%H:%M:%S', time.gmtime())}] {msg}", flush=True)


def node_rank_from_hostname(hostname: str) -> int:
    """StatefulSet convention: ``<base>-<ordinal>`` (reference ``entrypoint.sh:24-26``)."""
    tail = hostname.rsplit("-", 1)[-1]
    if not tail.isdigit():
        raise ValueError(f"hostname {hostname!r} has no numeric -<ordinal> suffix")
    return int(tail)


def master_from_hostname(hostname: str, discovery_service: str) -> str:
    """``<base>-0.<discovery service>`` (reference ``entrypoint.sh:26-28``)."""
    base = hostname.rsplit("-", 1)[0]
    return f"{base}-0.{discovery_service}"


def worker_env(base: Dict[str, str], *, rank: int, local_rank: int, world_size: int,
               nproc: int, node_rank: int, master_addr: str, master_port: int,
               run_id: str, restart_count: int = 0, max_restarts: int = 0) -> Dict[str, str]:
    env = dict(base)
    env.update({
        "RANK": str(rank),
        "LOCAL_RANK": str(local_rank),
        "WORLD_SIZE": str(world_size),
        "LOCAL_WORLD_SIZE": str(nproc),
        "GROUP_RANK": str(node_rank),
        "ROLE_RANK": str(rank),
        "ROLE_WORLD_SIZE": str(world_size),
        "GROUP_WORLD_SIZE": str(world_size // nproc),
        "MASTER_ADDR": master_addr,
        "MASTER_PORT": str(master_port),
        "TORCHELASTIC_RESTART_COUNT": str(restart_count),
        "TORCHELASTIC_MAX_RESTARTS": str(max_restarts),
        "TORCHELASTIC_RUN_ID": run_id,
        "TORCH_NCCL_ASYNC_ERROR_HANDLING": "1",
    })
    if nproc > 1 and "OMP_NUM_THREADS" not in base:
        env["OMP_NUM_THREADS"] = "1"
    return env


class Supervisor:
    def __init__(self, cmd: List[str], envs: List[Dict[str, str]]):
        self.cmd = cmd
        self.envs = envs
        self.procs: List[subprocess.Popen] = []
        self._stopping = False

    def start(self) -> None:
        for env in self.envs:
            self.procs.append(subprocess.Popen(self.cmd, env=env, start_new_session=True))

    def _signal_all(self, sig: int) -> None:
        for p in self.procs:
            if p.poll() is None:
                try:
                    os.killpg(p.pid, sig)
                except (ProcessLookupError, PermissionError):
                    pass

    def terminate(self) -> None:
        self._signal_all(signal.SIGTERM)
        deadline = time.time() + TERM_GRACE_S
        while time.time() < deadline and any(p.poll() is None for p in self.procs):
            time.sleep(0.05)
        self._signal_all(signal.SIGKILL)
        for p in self.procs:
            try:
                p.wait(timeout=5)
            except subprocess.TimeoutExpired:
                pass

    def on_signal(self, signum, _frame) -> None:
        self._stopping = True
        self._signal_all(signum)

    def wait(self) -> int:
        """Returns 0 when every worker exits 0; otherwise the first failure's code."""
        while True:
            running = 0
            for i, p in enumerate(self.procs):
                rc = p.poll()
                if rc is None:
                    running += 1
                elif rc != 0:
                    code = rc if rc > 0 else 128 - rc
                    log(f"ERROR: local rank {i} (pid {p.pid}) exited with code {rc}; "
                        f"stopping {running + sum(q.poll() is None for q in self.procs[i + 1:])} "
                        f"remaining worker(s)")
                    self.terminate()
                    return code
            if running == 0:
                return 0
            if self._stopping:
                self.terminate()
                return 130
            time.sleep(MONITOR_INTERVAL_S)


def build_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(prog="b200ddp.launch")
    p.add_argument("--nnodes", type=int, default=1)
    p.add_argument("--nproc-per-node", "--nproc_per_node", type=str, default="1",
                   help="worker count, or 'gpu' for one per visible GPU")
    p.add_argument("--node-rank", "--node_rank", type=int, default=0)
    p.add_argument("--master-addr", "--master_addr", type=str, default="127.0.0.1")
    p.add_argument("--master-port", "--master_port", type=int, default=29500)
    p.add_argument("--run-id", type=str, default=None)
    p.add_argument("--max-restarts", "--max_restarts", type=int, default=0,
                   help="restarts of the whole worker group after a failure (single node)")
    p.add_argument("--restart-args", type=str, default="",
                   help="extra training-script arguments on restarts, e.g. '--resume ck/latest_model.pt'")
    p.add_argument("--no-python", action="store_true", help="run the script as an executable")
    p.add_argument("-m", "--module", action="store_true", help="treat the script as a module name")
    p.add_argument("--dry-run", action="store_true", help="print the worker commands and exit")
    p.add_argument("training_script", type=str)
    p.add_argument("training_script_args", nargs=argparse.REMAINDER)
    return p


def resolve_nproc(spec: str) -> int:
    if spec == "gpu":
        try:
            out = subprocess.run(["nvidia-smi", "-L"], capture_output=True, text=True, timeout=20)
            n = sum(1 for line in out.stdout.splitlines() if line.startswith("GPU "))
        except (OSError, subprocess.TimeoutExpired):
            n = 0
        if n == 0:
            raise SystemExit("--nproc-per-node=gpu but no GPU is visible")
        return n
    return int(spec)


def main(argv: Optional[List[str]] = None) -> int:
    args = build_parser().parse_args(argv)
    nproc = resolve_nproc(args.nproc_per_node)
    world = args.nnodes * nproc
    if not (0 <= args.node_rank < args.nnodes):
        log(f"ERROR: node rank {args.node_rank} outside [0, {args.nnodes})")
        return 1
    run_id = args.run_id or os.environ.get("TORCHELASTIC_RUN_ID") or uuid.uuid4().hex[:12]
    if args.no_python:
        cmd = [args.training_script]
    elif args.module:
        cmd = [sys.executable, "-u", "-m", args.training_script]
    else:
        cmd = [sys.executable, "-u", args.training_script]
    cmd += args.training_script_args
    if args.max_restarts > 0 and args.nnodes > 1:
        log("ERROR: --max-restarts needs --nnodes 1 (static rendezvous cannot re-admit restarted nodes)")
        return 1

    def envs_for(attempt: int) -> List[Dict[str, str]]:
        return [
            worker_env(dict(os.environ), rank=args.node_rank * nproc + lr, local_rank=lr,
                       world_size=world, nproc=nproc, node_rank=args.node_rank,
                       master_addr=args.master_addr, master_port=args.master_port + attempt, run_id=run_id,
                       restart_count=attempt, max_restarts=args.max_restarts)
            for lr in range(nproc)
        ]

    if args.dry_run:
        for e in envs_for(0):
            print(f"RANK={e['RANK']} LOCAL_RANK={e['LOCAL_RANK']} WORLD_SIZE={e['WORLD_SIZE']} "
                  f"MASTER_ADDR={e['MASTER_ADDR']} MASTER_PORT={e['MASTER_PORT']} "
                  + " ".join(shlex.quote(c) for c in cmd))
        return 0
    log(f"node {args.node_rank}/{args.nnodes}: starting {nproc} worker(s), world size {world}, "
        f"rendezvous {args.master_addr}:{args.master_port} on {socket.gethostname()}")
    rc = 0
    for attempt in range(args.max_restarts + 1):
        attempt_cmd = cmd + (shlex.split(args.restart_args) if attempt > 0 else [])
        sup = Supervisor(attempt_cmd, envs_for(attempt))
        signal.signal(signal.SIGTERM, sup.on_signal)
        signal.signal(signal.SIGINT, sup.on_signal)
        sup.start()
        rc = sup.wait()
        if rc == 0 or sup._stopping or attempt == args.max_restarts:
            return rc
        log(f"worker group failed with code {rc}; restart {attempt + 1}/{args.max_restarts}"
            + (f" with {args.restart_args!r}" if args.restart_args else ""))
    return rc


if __name__ == "__main__":
    sys.exit(main())
