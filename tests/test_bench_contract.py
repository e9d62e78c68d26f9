"""bench.py's command-line contract (the driver calls it with --gpus/--steps/--warmup, --impl reference,
and with no flags at all) and the step-path description, checked without a GPU."""
import importlib.util
import os
import sys
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _load():
    spec = importlib.util.spec_from_file_location("bench_under_test", os.path.join(ROOT, "bench.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_defaults_finish_within_minutes_and_keep_three_warmup_steps(monkeypatch):
    bench = _load()
    monkeypatch.setattr(sys, "argv", ["bench.py"])
    a = bench.parse()
    assert a.gpus == 1 and a.impl == "b200ddp" and a.dtype == "bf16"
    assert a.warmup >= 3 and 20 <= a.steps <= 100000
    monkeypatch.setattr(sys, "argv", ["bench.py", "--impl", "reference", "--gpus", "8", "--steps", "20", "--warmup", "5"])
    a = bench.parse()
    assert (a.impl, a.gpus, a.steps, a.warmup) == ("reference", 8, 20, 5)


def test_step_path_names_what_ran():
    bench = _load()
    fused = types.SimpleNamespace(update_clusters=12)
    eng = types.SimpleNamespace(program=types.SimpleNamespace(fused=fused))
    assert "12 update clusters" in bench._step_path(eng) and "same launch" in bench._step_path(eng)
    fused.update_clusters = 0
    assert "fused cluster kernel" in bench._step_path(eng)
    assert "GEMM-per-layer" in bench._step_path(types.SimpleNamespace(program=types.SimpleNamespace(fused=None)))
    assert "torch autograd" in bench._step_path(types.SimpleNamespace(program=None))
