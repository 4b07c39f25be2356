"""echo_tts_b200 -- B200-native (sm_100a) Echo-TTS sampling engine behind the reference's Python API.

Importing the package does not touch the GPU; constructing a model or calling an op loads
``libecho_b200.so`` and fails loudly if it (or a CUDA device) is missing.
"""
from .schema import DiTConfig, AEConfig, ECHO_BASE, ECHO_TINY, AE_BASE, AE_TINY  # noqa: F401

__all__ = ["inference", "inference_blockwise", "model", "autoencoder", "sampler", "ops", "schema"]
