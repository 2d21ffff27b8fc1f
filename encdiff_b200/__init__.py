"""encdiff_b200 — B200-native (sm_100a) implementation of EncDiff's latent-diffusion denoising hot path."""
__version__ = "0.1.0"
