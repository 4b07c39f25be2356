"""Drop-in mirror of the reference ``inference.py`` public API (same names, argument meaning, error behaviour)
over the B200-native engine.  Host-side index logic (tokenizer, masks, length arithmetic) is plain Python and
bit-exact with the reference; everything numeric runs in the sm_100a kernels."""
from __future__ import annotations

from dataclasses import dataclass
from typing import Callable, List, Optional, Tuple

import torch

from . import ops
from .audio import load_audio, prepare_audio  # noqa: F401  (inference.py:141-149)
from .autoencoder import DAC
from .model import EchoDiT
from .sampler import sample_euler_cfg_independent_guidances  # noqa: F401  (re-export, inference.py:446)
from .schema import AE_BASE, ECHO_BASE, AEConfig, DiTConfig, ae_rope_table

KVCache = List[Tuple[torch.Tensor, torch.Tensor]]                       # inference.py:406
SampleFn = Callable[[EchoDiT, torch.Tensor, torch.Tensor, torch.Tensor, torch.Tensor, int], torch.Tensor]  # :341


@dataclass
class PCAState:                                                         # inference.py:116-121
    pca_components: torch.Tensor
    pca_mean: torch.Tensor
    latent_scale: float


# ---------------------------------------------------------------------------------------------
# loaders (inference.py:14-135) -- same signatures; weights are repacked for the kernels at load
# ---------------------------------------------------------------------------------------------

def load_model_from_state_dict(state: dict, device: str = "cuda", dtype=torch.bfloat16,
                               delete_blockwise_modules: bool = False, cfg: DiTConfig = ECHO_BASE) -> EchoDiT:
    if delete_blockwise_modules:                                        # inference.py:46-56
        state = {k: v for k, v in state.items()
                 if not (k.startswith("latent_encoder.") or k.startswith("latent_norm") or ".wk_latent" in k
                         or ".wv_latent" in k)}
    return EchoDiT(state, cfg=cfg, device=device, dtype=dtype if dtype is not None else torch.bfloat16)


def load_model_from_hf(repo_id: str = "jordand/echo-tts-base", device: str = "cuda",
                       dtype: Optional[torch.dtype] = torch.bfloat16, compile: bool = False,
                       token: Optional[str] = None, delete_blockwise_modules: bool = False) -> EchoDiT:
    """inference.py:14-69.  ``compile`` is accepted and ignored: there is no tracing compiler on this path."""
    from huggingface_hub import hf_hub_download
    import safetensors.torch as st
    w_path = hf_hub_download(repo_id, "pytorch_model.safetensors", token=token)
    state = st.load_file(w_path, device="cpu")
    return load_model_from_state_dict(state, device, dtype, delete_blockwise_modules)


def load_fish_ae_from_state_dict(state: dict, device: str = "cuda", dtype=torch.float32, cfg: AEConfig = AE_BASE) -> DAC:
    state = dict(state)
    # RoPE tables are module buffers (autoencoder.py:805-813); rebuild the ones a checkpoint does not carry
    enc_tr = f"encoder.block.{len(cfg.encoder_rates)}.block.5"
    for pm, args in (("quantizer.post_module", ()), ("quantizer.pre_module", ()), (enc_tr, (cfg.enc_tr_block_size, 64))):
        if f"{pm}.norm.weight" in state and f"{pm}.freqs_cis" not in state:
            state[f"{pm}.freqs_cis"] = ae_rope_table(cfg, *args)
    return DAC(state, cfg=cfg, device=device, dtype=dtype if dtype is not None else torch.float32)


def load_fish_ae_from_hf(repo_id: str = "jordand/fish-s1-dac-min", device: str = "cuda",
                         dtype: Optional[torch.dtype] = torch.float32, compile: bool = False,
                         token: Optional[str] = None) -> DAC:
    """inference.py:80-105."""
    from huggingface_hub import hf_hub_download
    import safetensors.torch as st
    w_path = hf_hub_download(repo_id, "pytorch_model.safetensors", token=token)
    return load_fish_ae_from_state_dict(st.load_file(w_path, device="cpu"), device, dtype)


def load_pca_state_from_hf(repo_id: str = "jordand/echo-tts-base", device: str = "cuda",
                           filename: str = "pca_state.safetensors", token: Optional[str] = None) -> PCAState:
    """inference.py:123-135."""
    from huggingface_hub import hf_hub_download
    import safetensors.torch as st
    t = st.load_file(hf_hub_download(repo_id, filename, token=token), device=device)
    return PCAState(pca_components=t["pca_components"], pca_mean=t["pca_mean"],
                    latent_scale=float(t["latent_scale"].item()))


def compile_model(model: EchoDiT) -> EchoDiT:                            # inference.py:72-77 (no-op here)
    return model


def compile_fish_ae(fish_ae: DAC) -> DAC:                                # inference.py:108-113 (no-op here)
    return fish_ae


# ---------------------------------------------------------------------------------------------
# tokenizer / masks (inference.py:152-217) -- host logic, bit-exact
# ---------------------------------------------------------------------------------------------

def tokenizer_encode(text: str, append_bos: bool = True, normalize: bool = True, return_normalized_text: bool = False):
    if normalize:
        text = (text.replace("…", "...").replace("’", "'").replace("”", '"').replace("\n", " ")
                .replace(":", ",").replace(";", ",").replace("—", ", "))
        if not text.startswith("[") and not text.startswith("(") and "S1" not in text and "S2" not in text:
            text = "[S1] " + text
    b = list(text.encode("utf-8"))
    if append_bos:
        b.insert(0, 0)
    if return_normalized_text:
        return torch.tensor(b), text
    return torch.tensor(b)


def get_text_input_ids_and_mask(text_arr: List[str], max_length: Optional[int], device: Optional[str] = None,
                                normalize: bool = True, return_normalized_text: bool = False, pad_to_max: bool = True):
    enc = [tokenizer_encode(t, normalize=normalize, return_normalized_text=True) for t in text_arr]
    if max_length is None:
        max_length = max(len(e) for e, _ in enc)
    tokens = torch.zeros((len(text_arr), max_length), dtype=torch.int32)
    mask = torch.zeros((len(text_arr), max_length), dtype=torch.bool)
    for i, (e, _) in enumerate(enc):
        n = min(len(e), max_length)
        tokens[i, :n] = e[:n]
        mask[i, :n] = 1
    if device is not None:
        tokens, mask = tokens.to(device), mask.to(device)
    if return_normalized_text:
        return tokens, mask, [t for _, t in enc]
    return tokens, mask


# ---------------------------------------------------------------------------------------------
# autoencoder glue (inference.py:223-309)
# ---------------------------------------------------------------------------------------------

@torch.inference_mode()
def ae_encode(fish_ae, pca_state: PCAState, audio: torch.Tensor) -> torch.Tensor:
    """inference.py:223-229.  The B200 DAC runs encoder + quantizer + PCA projection on its kernels
    (``DAC.encode_latent``); any other object with the reference's ``encode_zq`` is used duck-typed."""
    assert audio.ndim == 3 and audio.shape[1] == 1
    if isinstance(fish_ae, DAC):
        return fish_ae.encode_latent(audio, pca_state.pca_components, pca_state.pca_mean, pca_state.latent_scale)
    z_q = fish_ae.encode_zq(audio).float()
    z_q = (z_q.transpose(1, 2) - pca_state.pca_mean) @ pca_state.pca_components.T
    return z_q * pca_state.latent_scale


@torch.inference_mode()
def ae_decode(fish_ae: DAC, pca_state: PCAState, z_q: torch.Tensor) -> torch.Tensor:
    """inference.py:232-235 -> waveform fp32 [B, 1, T*2048]."""
    torch.cuda.nvtx.range_push("echo.ae_decode")
    wav = fish_ae.decode_latent(z_q, pca_state.pca_components, pca_state.pca_mean, pca_state.latent_scale).float()
    torch.cuda.nvtx.range_pop()
    return wav


@torch.inference_mode()
def ae_reconstruct(fish_ae, pca_state: PCAState, audio: torch.Tensor) -> torch.Tensor:
    assert audio.ndim == 3 and audio.shape[1] == 1
    return ae_decode(fish_ae, pca_state, ae_encode(fish_ae, pca_state, audio.to(fish_ae.dtype)))


@torch.inference_mode()
def get_speaker_latent_and_mask(fish_ae, pca_state: PCAState, audio: torch.Tensor, max_speaker_latent_length: int = 6400,
                                audio_chunk_size: int = 640 * 2048, pad_to_max: bool = False,
                                divis_by_patch_size: Optional[int] = 4, *, _encode_padding: bool = False
                                ) -> Tuple[torch.Tensor, torch.Tensor]:
    """inference.py:250-309 (chunked encode, mask, pad/trim, multiple-of-patch trim) -- length logic bit-exact.

    The reference pads the last chunk to 30 s of audio before encoding it (:268-271) and, unless ``pad_to_max``, throws the
    frames of the padding away again (:289-291).  The S1-DAC encoder is causal end to end (convolutions, window-limited
    transformers, downsampling), so the kept frames do not depend on the padding: without ``pad_to_max`` only the real
    samples (rounded up to a whole frame) are encoded -- a 10 s clip costs a third of the 30 s the reference encodes,
    with identical latents (``tests/test_gpu_encode.py::test_speaker_clip_trimmed_encode_equals_padded``)."""
    hop = 2048
    assert audio.ndim == 2 and audio.shape[0] == 1
    audio = audio[:, : max_speaker_latent_length * hop]
    chunks = []
    for i in range(0, audio.shape[1], audio_chunk_size):
        c = audio[:, i:i + audio_chunk_size]
        if c.shape[1] < audio_chunk_size:
            keep = audio_chunk_size if (pad_to_max or _encode_padding) else -(-c.shape[1] // hop) * hop
            c = torch.nn.functional.pad(c, (0, keep - c.shape[1]))
        chunks.append(ae_encode(fish_ae, pca_state, c.unsqueeze(0)))
    lat = torch.cat(chunks, dim=1)
    actual = audio.shape[1] // hop
    mask = (torch.arange(lat.shape[1], device=lat.device) < actual).unsqueeze(0)
    if pad_to_max and lat.shape[1] < max_speaker_latent_length:
        lat = torch.nn.functional.pad(lat, (0, 0, 0, max_speaker_latent_length - lat.shape[1]))
        mask = torch.nn.functional.pad(mask, (0, max_speaker_latent_length - mask.shape[1]))
    elif not pad_to_max:
        lat, mask = lat[:, :actual], mask[:, :actual]
    if divis_by_patch_size is not None:
        n = lat.shape[1] // divis_by_patch_size * divis_by_patch_size
        lat, mask = lat[:, :n], mask[:, :n]
    return lat, mask


# ---------------------------------------------------------------------------------------------
# flattening-point crop (inference.py:315-338): one kernel + one 4-byte read instead of <=640 x 2 syncs
# ---------------------------------------------------------------------------------------------

def find_flattening_point(data: torch.Tensor, target_value: float = 0.0, window_size: int = 20,
                          std_threshold: float = 0.05) -> int:
    idx = ops.flattening_point(data.float().contiguous(), window_size, std_threshold, target_value)
    return int(idx.item())


def crop_audio_to_flattening_point(audio: torch.Tensor, latent: torch.Tensor) -> torch.Tensor:
    return audio[..., : find_flattening_point(latent) * 2048]


@torch.inference_mode()
def sample_pipeline(model: EchoDiT, fish_ae, pca_state: PCAState, sample_fn: SampleFn, text_prompt: str,
                    speaker_audio: Optional[torch.Tensor], rng_seed: int,
                    pad_to_max_speaker_latent_length: Optional[int] = None,
                    pad_to_max_text_length: Optional[int] = None, normalize_text: bool = False) -> Tuple[torch.Tensor, str]:
    """inference.py:346-400.  ``speaker_audio`` may also be a ``(speaker_latent, speaker_mask)`` tuple of already
    encoded latents (extension: lets callers reuse a speaker they have already encoded)."""
    MAX_SPK, MAX_TEXT = 6400, 768
    device, dtype = model.device, model.dtype
    ids, mask, norm_text = get_text_input_ids_and_mask(
        [text_prompt], max_length=min(pad_to_max_text_length or MAX_TEXT, MAX_TEXT), device=device,
        normalize=normalize_text, return_normalized_text=True, pad_to_max=(pad_to_max_text_length is not None))
    if speaker_audio is None:
        spk = torch.zeros((1, pad_to_max_speaker_latent_length or 4, 80), device=device, dtype=dtype)
        smask = torch.zeros((1, pad_to_max_speaker_latent_length or 4), device=device, dtype=torch.bool)
    elif isinstance(speaker_audio, tuple):
        spk, smask = speaker_audio
    else:
        spk, smask = get_speaker_latent_and_mask(
            fish_ae, pca_state, speaker_audio.to(fish_ae.dtype).to(device),
            max_speaker_latent_length=pad_to_max_speaker_latent_length or MAX_SPK,
            pad_to_max=(pad_to_max_speaker_latent_length is not None))
    latent = sample_fn(model, spk, smask, ids, mask, rng_seed)
    # inference.py:395-397 decodes every latent and then cuts the waveform at the flattening point; the decoder is causal,
    # so the kept samples depend on the kept latents only: find the point first (it is a function of the latent) and
    # decode just that prefix
    crop = find_flattening_point(latent[0])
    if crop == 0:
        return torch.zeros((latent.shape[0], 1, 0), device=latent.device, dtype=torch.float32), norm_text[0]
    return ae_decode(fish_ae, pca_state, latent[:, :crop]), norm_text[0]
