"""Audio front-end of the sampling path: ``load_audio`` (reference ``inference.py:141-149``) and the
Gradio app's shape buckets (``gradio_app.py:139-166, 315-345``).

The reference decodes a file with torchcodec, averages the channels, resamples to 44.1 kHz with
``torchaudio.functional.resample`` and divides by ``max(peak, 1)`` on the CPU.  Here the file is decoded to PCM on the
host (torchcodec when it is installed, else the standard library for RIFF/WAVE), and everything after that -- channel
mean, polyphase windowed-sinc resampling, peak search, normalisation -- is two launches on the GPU
(``echo_resample_mono`` + ``echo_peak_normalize``), the result staying on the device for the S1-DAC encoder.
"""
from __future__ import annotations

import math
import struct
import wave
from typing import Dict, Optional, Tuple

import torch

from . import _lib as L

F32 = torch.float32
TARGET_SR = 44_100

_FILTER_BANKS: Dict[Tuple[int, int, str], tuple] = {}


def sinc_filter_bank(orig_freq: int, new_freq: int, lowpass_filter_width: int = 6, rolloff: float = 0.99):
    """The ``sinc_interp_hann`` filter bank of torchaudio 2.x ``_get_sinc_resample_kernel`` as
    ``functional.resample`` builds it for an fp32 waveform (every step in fp32): -> (bank fp32 [up, 2*width + down],
    width, down, up) with (down, up) = (orig, new) / gcd."""
    g = math.gcd(int(orig_freq), int(new_freq))
    down, up = int(orig_freq) // g, int(new_freq) // g
    base = min(down, up) * rolloff
    width = math.ceil(lowpass_filter_width * down / base)
    idx = torch.arange(-width, width + down, dtype=F32)[None] / down
    t = torch.arange(0, -up, -1, dtype=F32)[:, None] / up + idx
    t *= base
    t = t.clamp_(-lowpass_filter_width, lowpass_filter_width)
    window = torch.cos(t * math.pi / lowpass_filter_width / 2) ** 2
    t *= math.pi
    bank = torch.where(t == 0, torch.tensor(1.0), t.sin() / t)
    bank *= window * (base / down)
    return bank, width, down, up


def _device_bank(orig_freq: int, new_freq: int, device) -> tuple:
    key = (int(orig_freq), int(new_freq), str(device))
    ent = _FILTER_BANKS.get(key)
    if ent is None:
        if int(orig_freq) == int(new_freq):
            bank, width, down, up = torch.ones(1, 1), 0, 1, 1          # functional.resample returns the input unchanged
        else:
            bank, width, down, up = sinc_filter_bank(orig_freq, new_freq)
        # per-phase tap range that matters: beyond +-lowpass_filter_width zero crossings the Hann window is zero up to
        # rounding (|weight| < 1e-9 of the peak there; torchaudio multiplies through all 2*width + down taps regardless)
        live = bank.abs() > 1e-9 * float(bank.abs().max())
        idx = torch.arange(bank.shape[1])
        lo = torch.where(live, idx, bank.shape[1]).min(dim=1).values.to(torch.int32)
        hi = (torch.where(live, idx, -1).max(dim=1).values + 1).to(torch.int32)
        ent = (bank.t().contiguous().to(device), width, down, up, lo.to(device), hi.to(device))   # table [taps][up]
        _FILTER_BANKS[key] = ent
    return ent


@torch.inference_mode()
def prepare_audio(samples: torch.Tensor, sample_rate: int, device="cuda", target_rate: int = TARGET_SR,
                  normalize: bool = True) -> torch.Tensor:
    """``load_audio`` after the decode (inference.py:145-147): samples [channels, n] (any float/int dtype already scaled
    to [-1, 1], host or device) -> fp32 [1, ceil(n * target / rate)] on ``device``: channel mean, resample, / max(peak, 1)."""
    if not torch.cuda.is_available():
        raise RuntimeError("echo_tts_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
    if samples.dim() != 2 or samples.shape[1] == 0:
        raise ValueError("prepare_audio: samples must be [channels, n] with n > 0")
    x = samples.to(device, F32).contiguous()
    C_, n = x.shape
    table_t, width, down, up, tap_lo, tap_hi = _device_bank(sample_rate, target_rate, x.device)
    n_out = -(-up * n // down)
    y = torch.empty((1, n_out), device=x.device, dtype=F32)
    peak = torch.zeros(1, device=x.device, dtype=torch.int32)
    lib = L.load()
    L.check(lib.echo_resample_mono(x.data_ptr(), C_, n, x.stride(0), down, up, width, table_t.data_ptr(), tap_lo.data_ptr(),
                                   tap_hi.data_ptr(), y.data_ptr(), n_out, peak.data_ptr() if normalize else None, L.stream()))
    if normalize:
        L.check(lib.echo_peak_normalize(y.data_ptr(), n_out, peak.data_ptr(), 1.0, L.stream()))
    return y


def _decode_wav(path: str, max_duration: float) -> Tuple[torch.Tensor, int]:
    """RIFF/WAVE PCM (8/16/24/32-bit integer) -> (fp32 [channels, n] in [-1, 1), sample rate); standard library only."""
    with wave.open(path, "rb") as w:
        sr, ch, sw = w.getframerate(), w.getnchannels(), w.getsampwidth()
        n = min(w.getnframes(), int(max_duration * sr))
        raw = w.readframes(n)
    if sw == 1:
        a = (torch.frombuffer(bytearray(raw), dtype=torch.uint8).to(F32) - 128.0) / 128.0
    elif sw == 2:
        a = torch.frombuffer(bytearray(raw), dtype=torch.int16).to(F32) / 32768.0
    elif sw == 3:
        b = torch.frombuffer(bytearray(raw), dtype=torch.uint8).view(-1, 3).to(torch.int32)
        v = b[:, 0] | (b[:, 1] << 8) | (b[:, 2] << 16)
        a = (v - ((v & 0x800000) << 1)).to(F32) / 8388608.0
    elif sw == 4:
        a = torch.frombuffer(bytearray(raw), dtype=torch.int32).to(F32) / 2147483648.0
    else:
        raise ValueError(f"unsupported WAV sample width {sw}")
    return a.view(-1, ch).t().contiguous(), sr


def decode_file(path: str, max_duration: float = 300) -> Tuple[torch.Tensor, int]:
    """Host-side decode (inference.py:142-144): torchcodec when importable (any container / codec, as the reference),
    else PCM WAVE through the standard library.  -> (fp32 [channels, n], sample rate)."""
    try:
        from torchcodec.decoders import AudioDecoder
    except Exception:
        AudioDecoder = None
    if AudioDecoder is not None:
        dec = AudioDecoder(path)
        return dec.get_samples_played_in_range(0, max_duration).data.float(), int(dec.metadata.sample_rate)
    try:
        return _decode_wav(path, max_duration)
    except (wave.Error, EOFError, struct.error) as e:
        raise RuntimeError(f"{path}: torchcodec is not installed and the file is not PCM RIFF/WAVE ({e})") from e


def load_audio(path: str, max_duration: int = 300, device="cuda") -> torch.Tensor:
    """Drop-in for reference ``load_audio`` (inference.py:141-149): fp32 [1, n] at 44.1 kHz, peak <= 1 -- on the GPU
    (the reference returns a CPU tensor that every caller moves with ``.cuda()``, gradio_app.py:311-313)."""
    samples, sr = decode_file(path, max_duration)
    return prepare_audio(samples, sr, device=device)


# ---------------------------------------------------------------------------------------------------------------------
# shape buckets (gradio_app.py:139-166, 315-345): requests are padded to a few fixed shapes so that the captured CUDA
# graphs of the sampler (keyed by shape) are reused instead of re-captured
# ---------------------------------------------------------------------------------------------------------------------
def find_min_bucket_gte(values_str: str, actual_length: int) -> Optional[int]:
    """gradio_app.py:139-166: smallest comma-separated bucket >= actual_length (largest if none fits; a single value is
    returned as is; empty -> None)."""
    if not values_str or not values_str.strip():
        return None
    values_str = values_str.strip()
    if "," not in values_str:
        return int(values_str)
    values = [int(v.strip()) for v in values_str.split(",") if v.strip()]
    if not values:
        return None
    fits = [v for v in values if v >= actual_length]
    return min(fits) if fits else max(values)


def request_buckets(text_prompt: str, speaker_audio: Optional[torch.Tensor], text_buckets: str = "768",
                    speaker_buckets: str = "640,2816,6400") -> Tuple[Optional[int], Optional[int]]:
    """gradio_app.py:315-338: (pad_to_max_text_length, pad_to_max_speaker_latent_length) for ``sample_pipeline``."""
    text_len = len(text_prompt.encode("utf-8")) + 1                       # + BOS
    spk_len = (speaker_audio.shape[-1] // 2048) // 4 * 4 if speaker_audio is not None else 0
    return find_min_bucket_gte(text_buckets, text_len), find_min_bucket_gte(speaker_buckets, spk_len)
