"""Launcher: argv/env contract, error exits, failure propagation."""
import os
import stat
import subprocess
import sys
import textwrap
import time

import pytest

from b200ddp import launch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ENTRY = os.path.join(ROOT, "launch", "entrypoint.sh")


def _stub_bin(tmp_path, name, body):
    p = tmp_path / name
    p.write_text("#!/bin/bash\n" + body)
    p.chmod(p.stat().st_mode | stat.S_IEXEC)
    return p


def _run_entry(tmp_path, env, hostname="trainer-job-3"):
    _stub_bin(tmp_path, "hostname", f"echo {hostname}\n")
    _stub_bin(tmp_path, "torchrun", 'echo "ARGV:$*"\n')
    _stub_bin(tmp_path, "python", 'echo "ARGV:$*"\n')
    full = {"PATH": f"{tmp_path}:/usr/bin:/bin", **env}
    return subprocess.run(["bash", ENTRY], env=full, capture_output=True, text=True)


def test_entrypoint_argv_matches_reference_contract(tmp_path):
    r = _run_entry(tmp_path, {
        "NF_DISCOVERY_SERVICE": "disc.ns.svc", "REPLICAS": "4", "NPROC_PER_NODE": "8",
        "SCRIPT_ARGS": "--epochs 3 --batch-size 128", "LAUNCHER": "torchrun"})
    assert r.returncode == 0, r.stderr
    argv = [l for l in r.stdout.splitlines() if l.startswith("ARGV:")][0][5:]
    assert argv == ("--nnodes=4 --nproc-per-node=8 --node-rank=3 "
                    "--master-addr=trainer-job-0.disc.ns.svc --master-port=29500 "
                    "train.py --epochs 3 --batch-size 128")
    assert "Node 3/4 starting, master=trainer-job-0.disc.ns.svc" in r.stdout


def test_entrypoint_default_launcher_is_ours(tmp_path):
    r = _run_entry(tmp_path, {"NF_DISCOVERY_SERVICE": "d", "REPLICAS": "2"}, "job-1")
    argv = [l for l in r.stdout.splitlines() if l.startswith("ARGV:")][0][5:]
    assert argv == ("-m b200ddp.launch --nnodes=2 --nproc-per-node=1 --node-rank=1 "
                    "--master-addr=job-0.d --master-port=29500 -m b200ddp.train")


@pytest.mark.parametrize("env,msg", [
    ({"REPLICAS": "2"}, "ERROR: NF_DISCOVERY_SERVICE not set"),
    ({"NF_DISCOVERY_SERVICE": "d"}, "ERROR: REPLICAS not set"),
])
def test_entrypoint_hard_fails(tmp_path, env, msg):
    r = _run_entry(tmp_path, env)
    assert r.returncode == 1 and msg in r.stdout


def test_entrypoint_single_box_needs_no_discovery(tmp_path):
    r = _run_entry(tmp_path, {"REPLICAS": "1", "NPROC_PER_NODE": "8"}, "whatever")
    assert r.returncode == 0 and "--master-addr=127.0.0.1" in r.stdout
    assert "--node-rank=0" in r.stdout


def test_hostname_parsing():
    assert launch.node_rank_from_hostname("trainer-job-3") == 3
    assert launch.master_from_hostname("trainer-job-3", "svc") == "trainer-job-0.svc"
    with pytest.raises(ValueError):
        launch.node_rank_from_hostname("nohyphen")


def test_worker_env_contract():
    env = launch.worker_env({}, rank=11, local_rank=3, world_size=16, nproc=8, node_rank=1,
                            master_addr="m", master_port=29500, run_id="x")
    for k in ["RANK", "LOCAL_RANK", "WORLD_SIZE", "LOCAL_WORLD_SIZE", "GROUP_RANK", "ROLE_RANK",
              "MASTER_ADDR", "MASTER_PORT", "TORCHELASTIC_RESTART_COUNT",
              "TORCHELASTIC_MAX_RESTARTS", "TORCHELASTIC_RUN_ID",
              "TORCH_NCCL_ASYNC_ERROR_HANDLING"]:
        assert k in env
    assert env["RANK"] == "11" and env["GROUP_RANK"] == "1" and env["OMP_NUM_THREADS"] == "1"
    assert "OMP_NUM_THREADS" not in launch.worker_env(
        {}, rank=0, local_rank=0, world_size=1, nproc=1, node_rank=0, master_addr="m",
        master_port=1, run_id="x")


def test_supervisor_propagates_first_failure(tmp_path):
    script = tmp_path / "w.py"
    script.write_text(textwrap.dedent("""
        import os, sys, time
        if os.environ["RANK"] == "1":
            time.sleep(0.3); sys.exit(7)
        time.sleep(60)
    """))
    t0 = time.time()
    r = subprocess.run([sys.executable, "-m", "b200ddp.launch", "--nproc-per-node=3", str(script)],
                       cwd=ROOT, capture_output=True, text=True, timeout=60)
    assert r.returncode == 7
    assert time.time() - t0 < 20  # siblings were torn down, not waited for
    assert "exited with code 7" in r.stdout


def test_supervisor_success(tmp_path):
    script = tmp_path / "ok.py"
    script.write_text("import os; print('hello %s %s' % (os.environ['RANK'], os.environ['WORLD_SIZE']), flush=True)\n")
    r = subprocess.run([sys.executable, "-m", "b200ddp.launch", "--nproc-per-node=2", str(script)],
                       cwd=ROOT, capture_output=True, text=True, timeout=60)
    assert r.returncode == 0
    assert "hello 0 2" in r.stdout and "hello 1 2" in r.stdout


def test_max_restarts_relaunches_the_group_with_restart_args(tmp_path):
    """torchrun-agent semantics the reference would get with --max-restarts: the whole group is
    torn down and started again, TORCHELASTIC_RESTART_COUNT counts the attempt, and the extra
    arguments (a --resume in practice) reach the script only on restarts."""
    script = tmp_path / "flaky.py"
    script.write_text(textwrap.dedent("""
        import os, sys, time
        attempt = int(os.environ["TORCHELASTIC_RESTART_COUNT"])
        print("attempt %d rank %s port %s argv %s max %s" % (attempt, os.environ["RANK"], os.environ["MASTER_PORT"],
              sys.argv[1:], os.environ["TORCHELASTIC_MAX_RESTARTS"]), flush=True)
        if attempt == 0 and os.environ["RANK"] == "1":
            time.sleep(0.2); sys.exit(9)
        if attempt == 0:
            time.sleep(60)
    """))
    base = [sys.executable, "-m", "b200ddp.launch", "--nproc-per-node=2", "--master-port=29871"]
    r = subprocess.run(base + ["--max-restarts=1", "--restart-args=--resume ck.pt", str(script), "--epochs", "3"],
                       cwd=ROOT, capture_output=True, text=True, timeout=90)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "restart 1/1" in r.stdout
    assert "attempt 0 rank 0 port 29871 argv ['--epochs', '3'] max 1" in r.stdout
    assert "attempt 1 rank 1 port 29872 argv ['--epochs', '3', '--resume', 'ck.pt'] max 1" in r.stdout
    # default (the reference's configuration): no restart, the failure code comes back
    r0 = subprocess.run(base + [str(script)], cwd=ROOT, capture_output=True, text=True, timeout=90)
    assert r0.returncode == 9 and "attempt 1" not in r0.stdout


def test_max_restarts_is_single_node_only(tmp_path):
    r = subprocess.run([sys.executable, "-m", "b200ddp.launch", "--nnodes=2", "--max-restarts=1", "x.py"],
                       cwd=ROOT, capture_output=True, text=True, timeout=60)
    assert r.returncode == 1 and "--max-restarts needs --nnodes 1" in r.stdout
