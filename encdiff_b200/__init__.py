"""encdiff_b200 — B200-native (sm_100a) implementation of EncDiff's latent-diffusion denoising hot path.

Drop-in classes, selected with the ``target:`` lines of the reference's yaml: ``UNetModel`` (unet_config), ``Encoder4``
(cond_stage_config: concept encoder trunk + the per-unit MLPs, ``ConceptMLP``), ``VQModelInterface`` (first_stage_config: the
frozen VQ-f4 first stage; ``VQEncoder`` / ``VQDecoder`` are its two halves); ``SpatialTransformer``, ``CrossAttention``,
``DDIMSampler``; ``train.DenoiserTrainer`` / ``train.EncDiffTrainer`` are the fused training steps.
"""
__version__ = "0.1.0"

from .modules import CrossAttention, SpatialTransformer  # noqa: F401
from .unet import UNetModel  # noqa: F401
from .sampler import DDIMSampler  # noqa: F401
from .concept import ConceptMLP, Encoder4  # noqa: F401
from .first_stage import VQDecoder, VQEncoder, VQModelInterface  # noqa: F401
