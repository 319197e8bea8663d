"""Checkpoint contract, both directions through the unmodified reference helpers."""
import importlib.util
import os

import pytest
import torch

from mpi_pytorch_b200.engine.checkpoint import load_checkpoint, make_checkpoint, save_checkpoint
from mpi_pytorch_b200.models import initialize_model

REF = "/root/reference"


def _ref_module(name):
    path = os.path.join(REF, name + ".py")
    if not os.path.exists(path):
        pytest.skip("reference tree not mounted")
    spec = importlib.util.spec_from_file_location("ref_" + name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _train_a_bit(model, opt, steps=2):
    for _ in range(steps):
        loss = model(torch.randn(2, 3, 32, 32)).sum()
        opt.zero_grad()
        loss.backward()
        opt.step()


def test_layout_and_roundtrip(tmp_path):
    model, _ = initialize_model("resnet18", 7, False, use_pretrained=False)
    opt = torch.optim.Adam(model.parameters(), lr=4e-4)
    _train_a_bit(model, opt)
    ck = make_checkpoint(3, model, opt, 1.25)
    assert set(ck) == {"epoch", "state_dict", "optimizer", "loss"}
    path = save_checkpoint(ck, 3, "resnet18", str(tmp_path) + "/", str(tmp_path) + "/m/")
    assert os.path.basename(path) == "checkpoint_resnet18.pt"
    assert all(v.device.type == "cpu" for v in ck["state_dict"].values())
    m2, _ = initialize_model("resnet18", 7, False, use_pretrained=False)
    o2 = torch.optim.Adam(m2.parameters(), lr=4e-4)
    _, _, epoch = load_checkpoint(path, m2, o2)
    assert epoch == 3
    for a, b in zip(model.state_dict().values(), m2.state_dict().values()):
        assert torch.equal(a, b)


def test_ours_loads_in_reference_helpers(tmp_path):
    helpers = _ref_module("helpers")
    import torchvision
    model, _ = initialize_model("resnet18", 7, False, use_pretrained=False)
    opt = torch.optim.Adam(model.parameters(), lr=4e-4)
    _train_a_bit(model, opt)
    path = save_checkpoint(make_checkpoint(0, model, opt, 0.5), 0, "resnet18", str(tmp_path) + "/")
    ref_model = torchvision.models.resnet18(weights=None)
    ref_model.fc = torch.nn.Linear(512, 7)
    ref_opt = torch.optim.Adam(ref_model.parameters(), lr=4e-4)
    _orig = torch.load
    torch.load = lambda f, *a, **k: _orig(f, *a, **{**k, "weights_only": False})   # torch>=2.6 default differs
    try:
        m, o, e = helpers.load_checkpoint(path, ref_model, ref_opt)
    finally:
        torch.load = _orig
    assert e == 0
    x = torch.randn(1, 3, 32, 32)
    model.eval(), m.eval()
    assert torch.allclose(model(x), m(x), atol=1e-6)
    assert len(o.state_dict()["state"]) == len(opt.state_dict()["state"])


def test_reference_checkpoint_loads_in_ours(tmp_path):
    helpers = _ref_module("helpers")
    import torchvision
    ref_model = torchvision.models.resnet18(weights=None)
    ref_model.fc = torch.nn.Linear(512, 7)
    ref_opt = torch.optim.Adam(ref_model.parameters(), lr=4e-4)
    _train_a_bit(ref_model, ref_opt)
    state = {"epoch": 4, "state_dict": ref_model.state_dict(), "optimizer": ref_opt.state_dict(), "loss": 0.1}
    helpers.save_checkpoint(state, 4, "resnet18", str(tmp_path) + "/", str(tmp_path) + "/")
    model, _ = initialize_model("resnet18", 7, False, use_pretrained=False)
    opt = torch.optim.Adam(model.parameters(), lr=4e-4)
    _, _, epoch = load_checkpoint(str(tmp_path) + "/checkpoint_resnet18.pt", model, opt)
    assert epoch == 4
    x = torch.randn(1, 3, 32, 32)
    model.eval(), ref_model.eval()
    assert torch.allclose(model(x), ref_model(x), atol=1e-6)
