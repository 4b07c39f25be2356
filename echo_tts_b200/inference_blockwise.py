"""Mirror of the reference ``inference_blockwise.py`` public API, plus the streaming loop its README leaves out
("streaming ... not included", README.md:124; SURVEY 8 f2, BASELINE config 5)."""
from __future__ import annotations

from typing import Iterator, Optional, Sequence, Tuple

import torch

from .sampler import sample_blockwise_euler_cfg_independent_guidances, blockwise_blocks  # noqa: F401
from .streaming import DecodeStream, PrefixKVStream  # noqa: F401


@torch.inference_mode()
def stream_blockwise_audio(model, fish_ae, pca_state, speaker_latent: torch.Tensor, speaker_mask: torch.Tensor,
                           text_input_ids: torch.Tensor, text_mask: torch.Tensor, rng_seed: int,
                           block_sizes: Sequence[int], *, continuation_latent: Optional[torch.Tensor] = None,
                           decode_continuation: bool = False, **sampler_kwargs
                           ) -> Iterator[Tuple[torch.Tensor, int, int]]:
    """Blockwise generation (``inference_blockwise.py:14-123``) with each block decoded as soon as it is sampled:
    yields ``(audio fp32 [B, 1, (stop - start) * 2048], start, stop)`` per block.  The concatenated chunks equal
    ``ae_decode`` of the final latent (S1-DAC is causal: ``DecodeStream`` carries per-layer state), and the prefix KV
    grows incrementally (``PrefixKVStream``).  With a ``continuation_latent`` the decoder state is first advanced over
    it; its audio is yielded too when ``decode_continuation`` is set.  ``sampler_kwargs`` are the reference sampler's
    keyword arguments (num_steps, cfg_scale_text, ... speaker_kv_min_t)."""
    B = text_input_ids.shape[0]
    streams = fish_ae.stream_pool(B)            # kept across requests (state reset): their step graphs are reused

    def decode(lat: torch.Tensor) -> torch.Tensor:
        return torch.cat([st.step_latent(lat[b:b + 1], pca_state.pca_components, pca_state.pca_mean,
                                         pca_state.latent_scale) for b, st in enumerate(streams)], 0)

    if continuation_latent is not None and continuation_latent.shape[1] > 0:
        wav = decode(continuation_latent.to(model.device, torch.float32))
        if decode_continuation:
            yield wav, 0, continuation_latent.shape[1]
    for start, stop, prefix in blockwise_blocks(model, speaker_latent, speaker_mask, text_input_ids, text_mask, rng_seed,
                                                block_sizes, continuation_latent=continuation_latent, **sampler_kwargs):
        yield decode(prefix[:, start:stop]), start, stop
