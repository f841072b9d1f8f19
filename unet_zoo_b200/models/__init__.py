"""Model factory - the drop-in boundary of the hot path.

Mirrors the reference factory's contract (unet_zoo/models/__init__.py:59-61,78-238) for the four encoder-decoder
models that BASELINE.json names (plus `transatt_unet`, the first "next" model of SURVEY.md 8f rank 4): same signature, case-insensitive name lookup, the same always-popped keyword
arguments (``in_channels``, ``num_classes``, ``image_size``, ``depth``), the same per-model constructor arguments,
``ValueError`` for an unknown name and a warning (only) for ``pretrained=True``.  The reference's other nineteen
registry entries (transformer / axial-attention variants) are outside the hot path and raise NotImplementedError.
"""
from typing import Any, Dict, List

import torch.nn as nn

from .unet import UNet

_model_entries: Dict[str, Any] = {"unet": UNet}

try:  # the other three hot-path models register themselves as they are implemented
    from .attention_unet import AttentionUNet
    _model_entries["attention_unet"] = AttentionUNet
except ImportError:  # pragma: no cover
    pass
try:
    from .nested_unet import NestedUNet
    _model_entries["nested_unet"] = NestedUNet
except ImportError:  # pragma: no cover
    pass
try:
    from .resunet import ResUnet
    _model_entries["resunet"] = ResUnet
except ImportError:  # pragma: no cover
    pass

try:  # first neighbouring model on the same blocks
    from .transatt_unet import TransAttUNet
    _model_entries["transatt_unet"] = TransAttUNet
except ImportError:  # pragma: no cover
    pass

try:  # second neighbouring model: self- / cross-attention U-Net on the same blocks + the attention kernels
    from .unet_transformer import U_Transformer
    _model_entries["unet_transformer"] = U_Transformer
except ImportError:  # pragma: no cover
    pass

try:  # third neighbouring family: U^2-Net (RSU blocks = conv-BN-ReLU with dilation, nested encoder-decoders)
    from .u2net import U2NET, U2NETP
    _model_entries["u2net"] = U2NET
    _model_entries["u2netp"] = U2NETP
except ImportError:  # pragma: no cover
    pass

try:  # fourth: MultiResUnet (channel counts that are not multiples of 8 run zero-padded, models/_padded.py)
    from .multiresunet import MultiResUnet
    _model_entries["multiresunet"] = MultiResUnet
except ImportError:  # pragma: no cover
    pass

# registry names of the reference that are deliberately not part of this build
_OUT_OF_SCOPE = ("raunet", "da_transformer", "uctransnet",
                 "missformer", "vnet", "swin_unet_v2", "wranet", "egeunet", "unext", "unext_s",
                 "mmunet", "axialunet", "gated", "medt", "logo")


def list_models() -> List[str]:
    """Names accepted by :func:`create_model` (sorted, like the reference's list_models)."""
    return sorted(_model_entries)


def get_model_config(model_name: str, **kwargs) -> Dict[str, Any]:
    """The hot-path models have no default-config function in the reference (models/__init__.py:54-57,63-76)."""
    return {}


def create_model(model_name: str, pretrained: bool = False, **kwargs) -> nn.Module:
    name = model_name.lower()
    if name not in _model_entries:
        if name in _OUT_OF_SCOPE:
            raise NotImplementedError(
                f"'{model_name}' is a reference model outside the B200 hot path; available: {list_models()}")
        raise ValueError(f"Unknown model: '{model_name}'. Available models: {list_models()}")
    cls = _model_entries[name]

    in_channels = kwargs.pop("in_channels", 3)
    num_classes = kwargs.pop("num_classes", 1)
    kwargs.pop("image_size", None)   # accepted for every model, used by none of these four
    depth = kwargs.pop("depth", 5)
    precision = kwargs.pop("precision", "bf16")   # extension: "fp32" = forward-only parity mode (HotPathModel.set_precision)

    if name == "unet":
        args = dict(in_channels=in_channels, num_classes=num_classes)
    elif name == "attention_unet":
        args = dict(in_channels=in_channels, num_classes=num_classes, depth=depth)
    elif name == "nested_unet":
        args = dict(in_channels=in_channels, num_classes=num_classes, depth=depth, deep_supervision=False)
    elif name == "transatt_unet":    # models/__init__.py:106-109: `depth` is passed on and swallowed by **kwargs
        args = dict(in_channels=in_channels, num_classes=num_classes, depth=depth)
    elif name == "unet_transformer":   # the reference factory's default branch (models/__init__.py:221-223)
        args = dict(in_channels=in_channels, num_classes=num_classes)
    elif name == "multiresunet":       # models/__init__.py:134-137: `depth` is passed on and swallowed by **kwargs
        args = dict(in_channels=in_channels, num_classes=num_classes, depth=depth)
    elif name in ("u2net", "u2netp"):  # models/__init__.py:156-158
        args = dict(in_ch=in_channels, out_ch=num_classes)
    else:  # resunet
        args = dict(in_channels=in_channels, num_classes=num_classes, filters=[64, 128, 256, 512])
    args.update(kwargs)              # leftovers reach the constructor (TypeError if it does not know them)
    model = cls(**args)
    model.set_precision(precision)
    if pretrained:
        print(f"Warning: Pretrained weights not implemented for {model_name}. Initializing with random weights.")
    return model
