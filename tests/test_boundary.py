"""Drop-in boundary on CPU: factory contract, state_dict layout, error behaviour, C-ABI symbol table."""
import ctypes
import os
import re

import pytest
import torch

from conftest import ROOT, load_golden, seeded_model
import unet_zoo_b200
from unet_zoo_b200 import _lib
from unet_zoo_b200.models import blocks


def test_list_models_and_names():
    assert unet_zoo_b200.list_models() == ["attention_unet", "multiresunet", "nested_unet", "resunet", "transatt_unet",
                                           "u2net", "u2netp", "unet", "unet_transformer"]
    names = {"unet": "UNet", "attention_unet": "AttentionUNet", "nested_unet": "NestedUNet", "resunet": "ResUnet",
             "transatt_unet": "TransAttUNet", "unet_transformer": "U_Transformer",
             "u2net": "U2NET", "u2netp": "U2NETP", "multiresunet": "MultiResUnet"}
    for n, cls in names.items():
        assert type(unet_zoo_b200.create_model(n.upper())).__name__ == cls      # case-insensitive (__init__.py:82)


@pytest.mark.parametrize("name,count", [("unet", 136), ("attention_unet", 240), ("nested_unet", 212), ("resunet", 127),
                                        ("transatt_unet", 137), ("unet_transformer", 209),
                                        ("u2net", 798), ("u2netp", 798), ("multiresunet", 350)])
def test_default_init_matches_reference(name, count):
    """Same keys, shapes and - under the same seed - bit-identical default initialisation as the reference modules."""
    fp = load_golden(name)["fingerprint"]
    sd = seeded_model(name).state_dict()
    assert len(sd) == count and list(sd) == list(fp)
    for k, v in sd.items():
        assert float(v.double().sum()) == fp[k][0] and float(v.double().abs().sum()) == fp[k][1], k


def test_factory_kwargs_contract(capsys):
    m = unet_zoo_b200.create_model("unet", in_channels=1, num_classes=3, image_size=512, depth=4)   # popped silently
    assert m.down_convolution_1.conv.conv_op[0].in_channels == 1 and m.out.conv.out_channels == 3
    with pytest.raises(ValueError, match="Unknown model"):
        unet_zoo_b200.create_model("not_a_model")
    with pytest.raises(NotImplementedError):
        unet_zoo_b200.create_model("swin_unet_v2")
    with pytest.raises(TypeError):
        unet_zoo_b200.create_model("unet", bogus=1)            # leftovers reach the constructor (__init__.py:228-233)
    unet_zoo_b200.create_model("unet", pretrained=True)
    assert "Pretrained weights not implemented" in capsys.readouterr().out
    ds = unet_zoo_b200.create_model("nested_unet", deep_supervision=True)
    assert hasattr(ds, "final4") and not hasattr(ds, "final")
    t = unet_zoo_b200.create_model("transatt_unet", depth=5, image_size=512)       # both swallowed, as in the reference
    assert t.sdpa.dropout.p == 0.1 and t.pos.row_embed.weight.shape == (32, 256)
    with pytest.raises(NotImplementedError, match="bilinear"):
        unet_zoo_b200.create_model("transatt_unet", bilinear=False)
    r = unet_zoo_b200.create_model("resunet", filters=[32, 64, 128, 256], num_classes=2)
    assert r.output_layer[0].out_channels == 2 and "Warning: ResUnet" in capsys.readouterr().out


def test_no_cpu_fallback():
    m = unet_zoo_b200.create_model("unet")
    with pytest.raises(RuntimeError, match="no CPU path"):
        m(torch.randn(1, 3, 64, 64))
    with pytest.raises(RuntimeError, match="no stand-alone eager path"):
        m.bottle_neck(torch.randn(1, 512, 4, 4))
    with pytest.raises(RuntimeError):
        blocks.DoubleConv(3, 8)(torch.randn(1, 3, 8, 8))


def test_state_dict_roundtrip_with_module_prefix(tmp_path):
    """Reference checkpoints (multi_gpu.py:39-53: plain state_dict, optional 'module.' prefix) load unchanged."""
    a, b = seeded_model("unet"), unet_zoo_b200.create_model("unet")
    path = tmp_path / "unet_last.pth"
    torch.save({"module." + k: v for k, v in a.state_dict().items()}, path)
    sd = torch.load(path)
    sd = {k[len("module."):]: v for k, v in sd.items()}
    b.load_state_dict(sd)
    for (k, v), (_, w) in zip(a.state_dict().items(), b.state_dict().items()):
        assert torch.equal(v, w), k


def _header_symbols():
    text = open(os.path.join(ROOT, "include", "unet_zoo_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(uz_[a-z0-9_]+)\s*\(", text)))


def test_header_and_binding_agree():
    assert _header_symbols() == _lib.exported_symbols()


def test_library_exports_every_declared_symbol():
    if not os.path.exists(_lib.LIB_PATH):
        import __graft_entry__
        __graft_entry__.build()
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for sym in _header_symbols():
        assert hasattr(lib, sym), sym
    assert lib.uz_version() >= 100
    # no device here: the library must report that instead of computing anything on the host
    if not torch.cuda.is_available():
        lib.uz_last_error.restype = ctypes.c_char_p
        assert lib.uz_device_sm_count() < 0 and lib.uz_last_error()


def test_precision_and_input_normalization_knobs_are_validated_on_the_host():
    """Extensions of the factory / module surface: `precision` ("bf16" default, "fp32" forward-only parity mode) and
    `set_input_normalization` (decoded uint8 batches).  Bad values fail on the host, before any device work."""
    m = unet_zoo_b200.create_model("unet", precision="fp32")
    assert m.precision == "fp32" and unet_zoo_b200.create_model("unet").precision == "bf16"
    with pytest.raises(ValueError):
        unet_zoo_b200.create_model("unet", precision="fp16")
    with pytest.raises(ValueError):
        m.set_input_normalization((0.5,), (0.5,))               # needs one entry per input channel
    with pytest.raises(ValueError):
        m.set_input_normalization((0.5, 0.5, 0.5), (0.5, 0.0, 0.5))
    assert m.set_input_normalization((0.5, 0.5, 0.5), (0.2, 0.2, 0.2)) is m
    assert len(m.state_dict()) == 136                           # neither knob adds parameters or buffers
    with pytest.raises(RuntimeError):                           # still no CPU path, in either precision
        m(torch.zeros(1, 3, 32, 32))
