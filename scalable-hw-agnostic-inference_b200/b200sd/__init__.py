"""b200sd -- B200-native (sm_100a) Stable Diffusion 2.1 denoising hot path behind the reference's plugin seam.

Importing this package loads libb200sd.so; there is no CPU or PyTorch fallback for any compute op.
"""
from . import _lib

_lib.load()  # fail loudly if the CUDA library is missing

from . import ops  # noqa: E402
from .clip import B200CLIPTextEncoder, CLIPTextConfig, HashTokenizer, PromptEncoder  # noqa: E402
from .config import UNetConfig, VAEConfig  # noqa: E402
from .pipeline import B200Pipeline, StandInTextEncoder, install_into  # noqa: E402
from .scheduler import B200Scheduler  # noqa: E402
from .unet import B200UNet  # noqa: E402
from .vae import B200PostQuantConv, B200QuantConv, B200VAEDecoder, B200VAEEncoder  # noqa: E402

__all__ = ["ops", "B200CLIPTextEncoder", "CLIPTextConfig", "HashTokenizer", "PromptEncoder", "UNetConfig", "VAEConfig", "B200Pipeline", "StandInTextEncoder", "install_into", "B200Scheduler",
           "B200UNet", "B200VAEDecoder", "B200PostQuantConv", "B200VAEEncoder", "B200QuantConv"]
