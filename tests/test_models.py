"""Model zoo parity with torchvision (state_dict names/shapes, forward equality) -- SURVEY C9/C10."""
import warnings

import pytest
import torch
import torchvision

from mpi_pytorch_b200.models import MODEL_NAMES, initialize_model

TV = {"resnet18": torchvision.models.resnet18, "resnet34": torchvision.models.resnet34,
      "alexnet": torchvision.models.alexnet, "vgg": torchvision.models.vgg11_bn, "vgg16": torchvision.models.vgg16,
      "squeezenet": torchvision.models.squeezenet1_0, "densenet": torchvision.models.densenet121,
      "inception": torchvision.models.inception_v3}
INPUT = {"resnet18": 224, "resnet34": 128, "alexnet": 224, "vgg": 224, "vgg16": 224, "squeezenet": 224,
         "densenet": 224, "inception": 299}


@pytest.mark.parametrize("name", MODEL_NAMES)
def test_state_dict_matches_torchvision(name):
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        m, size = initialize_model(name, 1000, False, use_pretrained=False)
        kw = dict(weights=None)
        if name == "inception":
            kw.update(aux_logits=True, init_weights=False)
        ref = TV[name](**kw)
    assert size == INPUT[name]
    a = [(k, tuple(v.shape)) for k, v in m.state_dict().items()]
    b = [(k, tuple(v.shape)) for k, v in ref.state_dict().items()]
    assert a == b


@pytest.mark.parametrize("name", ["resnet18", "squeezenet", "densenet"])
def test_forward_equals_torchvision(name):
    m, _ = initialize_model(name, 1000, False, use_pretrained=False)
    ref = TV[name](weights=None)
    ref.load_state_dict(m.state_dict())
    m.eval(), ref.eval()
    x = torch.randn(2, 3, 64, 64)
    with torch.no_grad():
        assert torch.allclose(m(x), ref(x), atol=1e-5)


def test_head_replaced_and_freeze():
    m, _ = initialize_model("resnet18", 17, True, use_pretrained=False)
    assert m.fc.out_features == 17
    trainable = [n for n, p in m.named_parameters() if p.requires_grad]
    assert trainable == ["fc.weight", "fc.bias"]                      # models.py:5-13 semantics
    m, _ = initialize_model("squeezenet", 9, False, use_pretrained=False)
    assert m.classifier[1].out_channels == 9 and m.num_classes == 9
    m, _ = initialize_model("inception", 11, False, use_pretrained=False)
    assert m.fc.out_features == 11 and m.AuxLogits.fc.out_features == 11


def test_invalid_name_raises():
    with pytest.raises(ValueError):
        initialize_model("resnet50", 10, False, use_pretrained=False)


def test_inception_training_outputs_and_small_input():
    m, _ = initialize_model("inception", 5, False, use_pretrained=False)
    m.train()
    out = m(torch.randn(2, 3, 128, 128))          # aux head cannot run at 128x128 -> None instead of a crash (D2)
    assert out.logits.shape == (2, 5) and out.aux_logits is None
