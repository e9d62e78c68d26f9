"""Checkpoint compatibility with the reference in both directions."""
import copy
import zipfile

import torch

from b200ddp.checkpoint import load_checkpoint, save_checkpoint
from b200ddp.models.simplenet import SimpleNet
from b200ddp.parallel.engine import DataParallel

KEYS = ["layers.0.weight", "layers.0.bias", "layers.3.weight", "layers.3.bias",
        "layers.6.weight", "layers.6.bias"]


def _trained_engine(steps=3, seed=0):
    torch.manual_seed(seed)
    eng = DataParallel(SimpleNet(), optimizer="adam", lr=1e-3, backend="gloo")
    for _ in range(steps):
        eng.train_step(torch.randn(16, 784), torch.randint(0, 10, (16,)))
    return eng


def test_file_layout(tmp_path):
    eng = _trained_engine()
    path = str(tmp_path / "latest_model.pt")
    save_checkpoint(eng, eng.optimizer, 1, 2.25, path)
    ck = torch.load(path, map_location="cpu", weights_only=False)
    assert set(ck) == {"epoch", "model_state_dict", "optimizer_state_dict", "loss"}
    assert isinstance(ck["epoch"], int) and isinstance(ck["loss"], float)
    assert list(ck["model_state_dict"]) == KEYS
    for v in ck["model_state_dict"].values():
        assert v.dtype == torch.float32 and v.is_contiguous()
        assert v.untyped_storage().nbytes() == v.numel() * 4  # own storage, not the arena
    st = ck["optimizer_state_dict"]["state"]
    assert sorted(st) == [0, 1, 2, 3, 4, 5]
    assert st[0]["step"].dim() == 0 and st[0]["step"].dtype == torch.float32
    assert float(st[0]["step"]) == 3.0
    with zipfile.ZipFile(path) as z:
        storages = [n for n in z.namelist() if "/data/" in n and not n.endswith("serialization_id")]
    assert len(storages) == 24  # 6 params + 6 x (step, exp_avg, exp_avg_sq)


def test_reference_loads_our_file(tmp_path, reference_module):
    ref = reference_module
    eng = _trained_engine()
    path = str(tmp_path / "ours.pt")
    save_checkpoint(eng, eng.optimizer, 4, 1.5, path)
    model = ref.SimpleNet()
    opt = torch.optim.Adam(model.parameters(), lr=0.5)
    epoch = ref.load_checkpoint(model, opt, path, torch.device("cpu"))
    assert epoch == 4
    for (n, p), q in zip(model.named_parameters(), eng.module.parameters()):
        assert torch.equal(p, q), n
    # one more identical step on both sides stays bit-identical
    model.eval(); eng.eval()
    x = torch.randn(16, 784); y = torch.randint(0, 10, (16,))
    opt.zero_grad(); torch.nn.functional.cross_entropy(model(x), y).backward(); opt.step()
    eng.zero_grad(); eng.begin_backward()
    torch.nn.functional.cross_entropy(eng.module(x), y).backward(); eng.finish_step()
    for (n, p), q in zip(model.named_parameters(), eng.module.parameters()):
        assert torch.equal(p, q), n


def test_we_load_reference_file(tmp_path, reference_module):
    ref = reference_module
    torch.manual_seed(3)
    model = ref.SimpleNet()
    opt = torch.optim.Adam(model.parameters(), lr=2e-3)
    for _ in range(4):
        opt.zero_grad()
        torch.nn.functional.cross_entropy(model(torch.randn(8, 784)),
                                          torch.randint(0, 10, (8,))).backward()
        opt.step()
    path = str(tmp_path / "ref.pt")
    ref.save_checkpoint(model, opt, 7, 0.75, path)
    eng = DataParallel(SimpleNet(), optimizer="adam", lr=1e-3, backend="gloo")
    assert load_checkpoint(eng, eng.optimizer, path, torch.device("cpu")) == 7
    assert eng.optimizer.step_count == 4 and eng.optimizer.hp["lr"] == 2e-3
    for (n, p), q in zip(model.named_parameters(), eng.module.parameters()):
        assert torch.equal(p, q), n
    model.eval(); eng.eval()
    x = torch.randn(16, 784); y = torch.randint(0, 10, (16,))
    opt.zero_grad(); torch.nn.functional.cross_entropy(model(x), y).backward(); opt.step()
    eng.zero_grad(); eng.begin_backward()
    torch.nn.functional.cross_entropy(eng.module(x), y).backward(); eng.finish_step()
    for (n, p), q in zip(model.named_parameters(), eng.module.parameters()):
        assert torch.equal(p, q), n


def test_reference_save_accepts_our_objects(tmp_path, reference_module):
    """The engine quacks like DDP (.module) and the optimizer like torch.optim."""
    eng = _trained_engine()
    path = str(tmp_path / "via_ref.pt")
    reference_module.save_checkpoint(eng, eng.optimizer, 2, 1.0, path)
    eng2 = DataParallel(SimpleNet(), optimizer="adam", backend="gloo")
    assert load_checkpoint(eng2, eng2.optimizer, path, torch.device("cpu")) == 2
    assert torch.equal(eng2.master, eng.master)


def test_atomic_write_leaves_no_temp(tmp_path):
    eng = _trained_engine(1)
    save_checkpoint(eng, eng.optimizer, 0, 1.0, str(tmp_path / "a.pt"))
    assert sorted(p.name for p in tmp_path.iterdir()) == ["a.pt"]
