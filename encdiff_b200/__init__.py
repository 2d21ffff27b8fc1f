"""encdiff_b200 — B200-native (sm_100a) implementation of EncDiff's latent-diffusion denoising hot path.

Drop-in classes (select with ``target: encdiff_b200.UNetModel`` in the reference's unet_config):
``UNetModel``, ``SpatialTransformer``, ``CrossAttention``, ``DDIMSampler``; ``ConceptMLP`` replaces the per-unit MLP loop
of the reference's ``Encoder4`` (``net`` / ``warp``); ``VQEncoder`` / ``VQDecoder`` are the encode / decode halves of the frozen VQ-f4 first stage.
"""
__version__ = "0.1.0"

from .modules import CrossAttention, SpatialTransformer  # noqa: F401
from .unet import UNetModel  # noqa: F401
from .sampler import DDIMSampler  # noqa: F401
from .concept import ConceptMLP, Encoder4  # noqa: F401
from .first_stage import VQDecoder, VQEncoder  # noqa: F401
