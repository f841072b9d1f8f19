"""unet_zoo_b200 - B200-native (sm_100a) implementation of the UNet encoder-decoder hot path of unet_zoo."""
from .models import create_model, get_model_config, list_models  # noqa: F401

__version__ = "0.1.0"
