"""Mirror of the reference ``inference_blockwise.py`` public API."""
from .sampler import sample_blockwise_euler_cfg_independent_guidances  # noqa: F401
