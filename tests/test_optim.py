"""Flat optimizers vs torch.optim — bit-exact in fp32."""
import copy

import pytest
import torch

from b200ddp.models.simplenet import SimpleNet
from b200ddp.parallel.engine import DataParallel


def _pair(kind, **kw):
    torch.manual_seed(0)
    ref_model = SimpleNet()
    our_model = copy.deepcopy(ref_model)
    eng = DataParallel(our_model, optimizer=kind, backend="gloo", **kw)
    cls = {"adam": torch.optim.Adam, "adamw": torch.optim.AdamW, "sgd": torch.optim.SGD}[kind]
    ref_opt = cls(ref_model.parameters(), **kw)
    return ref_model, ref_opt, eng


@pytest.mark.parametrize("kind,kw", [
    ("adam", dict(lr=1e-3)),
    ("adam", dict(lr=3e-3, weight_decay=0.1)),
    ("adamw", dict(lr=1e-3, weight_decay=0.05)),
    ("sgd", dict(lr=0.05, momentum=0.9)),
    ("sgd", dict(lr=0.05, momentum=0.9, weight_decay=1e-3)),
])
def test_bit_exact_vs_torch(kind, kw):
    ref_model, ref_opt, eng = _pair(kind, **kw)
    ref_model.eval(); eng.eval()  # dropout off so both see identical graphs
    torch.manual_seed(1)
    for _ in range(5):
        x = torch.randn(32, 784); y = torch.randint(0, 10, (32,))
        ref_opt.zero_grad()
        torch.nn.functional.cross_entropy(ref_model(x), y).backward()
        ref_opt.step()
        eng.zero_grad(); eng.begin_backward()
        torch.nn.functional.cross_entropy(eng.module(x), y).backward()
        eng.finish_step()
    for (n, p), q in zip(ref_model.named_parameters(), eng.module.parameters()):
        assert torch.equal(p, q), n


def test_state_dict_layout_matches_torch_adam():
    ref_model, ref_opt, eng = _pair("adam", lr=1e-3)
    x = torch.randn(8, 784); y = torch.randint(0, 10, (8,))
    torch.nn.functional.cross_entropy(ref_model(x), y).backward(); ref_opt.step()
    eng.train_step(x, y)
    a, b = ref_opt.state_dict(), eng.optimizer.state_dict()
    assert a["param_groups"][0].keys() == b["param_groups"][0].keys()
    assert a["param_groups"][0]["params"] == b["param_groups"][0]["params"]
    assert a["state"].keys() == b["state"].keys()
    for k in a["state"]:
        assert a["state"][k].keys() == b["state"][k].keys()
        for name in a["state"][k]:
            ta, tb = a["state"][k][name], b["state"][k][name]
            assert ta.shape == tb.shape and ta.dtype == tb.dtype and ta.device == tb.device
    assert float(b["state"][0]["step"]) == 1.0


def test_wire_dtype_other_than_parameter_dtype_still_trains():
    """fp32 compute with bf16 gradients on the wire: autograd hands each gradient over as a fresh
    fp32 tensor, the hook casts it into the bf16 arena the reduction reads (round-1 finding: the
    arena stayed zero and the parameters never moved)."""
    torch.manual_seed(0)
    eng = DataParallel(SimpleNet(dropout=0.0), backend="gloo", compute_dtype=torch.float32,
                       wire_dtype=torch.bfloat16, lr=1e-3)
    ref = copy.deepcopy(eng.module)
    before = eng.master.clone()
    x = torch.randn(16, 784); y = torch.randint(0, 10, (16,))
    eng.train_step(x, y)
    torch.nn.functional.cross_entropy(ref(x), y).backward()
    for p, slot in zip(ref.parameters(), eng.layout.slots):
        got = eng.grad[slot.offset:slot.offset + slot.numel].float().view(slot.shape)
        assert torch.allclose(got, p.grad, rtol=1e-2, atol=1e-4), slot.name
    assert (eng.master - before).abs().max().item() > 1e-4
    assert all(p.grad is None for p in eng.module.parameters())      # moved, not kept twice


def test_backend_auto_picks_by_topology(monkeypatch):
    from b200ddp.parallel import process_group as pg
    assert pg.pick_engine_backend("nccl", "cuda:0") == "nccl"
    assert pg.pick_engine_backend("auto", "cpu") == "gloo"
    assert pg.pick_engine_backend("auto", "cuda:0") == "b200"          # world of one
    # two ranks on two hosts: peer-mapped arenas are impossible, NCCL carries the gradients
    monkeypatch.setattr(pg.dist, "is_initialized", lambda: True)
    monkeypatch.setattr(pg.dist, "get_world_size", lambda: 2)
    def fake_gather(out, obj):
        out[0], out[1] = "node-0", "node-1"
    monkeypatch.setattr(pg.dist, "all_gather_object", fake_gather)
    assert pg.pick_engine_backend("auto", "cuda:0") == "nccl"
    def same_host(out, obj):
        out[0] = out[1] = "node-0"
    monkeypatch.setattr(pg.dist, "all_gather_object", same_host)
    monkeypatch.setenv("LOCAL_WORLD_SIZE", "2")
    assert pg.pick_engine_backend("auto", "cuda:0") == "b200"
    monkeypatch.setattr(pg.dist, "get_world_size", lambda: 16)
    def sixteen(out, obj):
        for i in range(16):
            out[i] = "node-0"
    monkeypatch.setattr(pg.dist, "all_gather_object", sixteen)
    monkeypatch.setenv("LOCAL_WORLD_SIZE", "16")
    assert pg.pick_engine_backend("auto", "cuda:0") == "nccl"
