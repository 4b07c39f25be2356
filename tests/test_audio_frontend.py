"""Audio front-end (load_audio, inference.py:141-149).  CPU part: the oracle's restatement of the third-party
resampler against the installed torchaudio, the host-built filter bank, the WAV decode and the Gradio bucket logic
(against the reference's own function, extracted from gradio_app.py without importing gradio).  GPU part: the
resample / normalise kernels against the oracle."""
import ast
import math
import os
import struct
import wave

import pytest
import torch

from conftest import rel_l2

RATES = [(48_000, 44_100), (16_000, 44_100), (22_050, 44_100), (96_000, 44_100), (8_000, 44_100), (44_100, 44_100)]


def _sig(channels, n, seed=0):
    g = torch.Generator().manual_seed(seed)
    t = torch.arange(n) / 1000.0
    return 0.6 * torch.sin(2 * math.pi * 0.7 * t)[None] + 0.5 * torch.randn(channels, n, generator=g)


@pytest.mark.parametrize("orig,new", RATES)
def test_oracle_resampler_matches_torchaudio(orig, new):
    ta = pytest.importorskip("torchaudio")
    from oracle import echo_oracle as O
    x = _sig(1, 7001)
    ref = ta.functional.resample(x, orig, new)
    got = O.resample_sinc(x, orig, new)
    assert got.shape == ref.shape
    assert torch.equal(got, ref)                       # same ops in the same order: bit-exact


@pytest.mark.parametrize("orig,new", RATES[:-1])
def test_filter_bank_matches_torchaudio(orig, new):
    ta = pytest.importorskip("torchaudio")
    from torchaudio.functional.functional import _get_sinc_resample_kernel
    from echo_tts_b200.audio import sinc_filter_bank
    g = math.gcd(orig, new)
    ref, width = _get_sinc_resample_kernel(orig, new, g, dtype=torch.float32)
    bank, w, down, up = sinc_filter_bank(orig, new)
    assert (w, down, up) == (width, orig // g, new // g)
    assert torch.equal(bank, ref[:, 0])


def test_load_audio_post_live_against_reference_lines():
    """inference.py:145-147 executed literally with torchaudio vs the oracle."""
    ta = pytest.importorskip("torchaudio")
    from oracle import echo_oracle as O
    x = 1.7 * _sig(2, 30_000, seed=3)
    audio = x.mean(dim=0).unsqueeze(0)
    audio = ta.functional.resample(audio, 48_000, 44_100)
    audio = audio / torch.maximum(audio.abs().max(), torch.tensor(1.0))
    assert torch.equal(O.load_audio_post(x, 48_000), audio)
    assert float(audio.abs().max()) <= 1.0


def test_bucket_logic_matches_reference_function():
    from echo_tts_b200.audio import find_min_bucket_gte, request_buckets
    path = "/root/reference/gradio_app.py"
    cases = [("", 5), ("  ", 5), ("768", 900), ("640, 2816, 6400", 0), ("640, 2816, 6400", 640), ("640, 2816, 6400", 641),
             ("640,2816,6400", 7000), ("6400, 640", 100), (",", 3), ("12,", 13)]
    if os.path.isfile(path):
        tree = ast.parse(open(path).read())
        fn = next(n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "find_min_bucket_gte")
        ns = {}
        exec(compile(ast.Module([fn], []), path, "exec"), ns)
        for s, n in cases:
            assert find_min_bucket_gte(s, n) == ns["find_min_bucket_gte"](s, n), (s, n)
    want = [None, None, 768, 640, 640, 2816, 6400, 640, None, 12]
    assert [find_min_bucket_gte(s, n) for s, n in cases] == want
    assert request_buckets("[S1] hi", torch.zeros(1, 10 * 44100)) == (768, 640)
    assert request_buckets("x" * 900, torch.zeros(1, 700 * 2048)) == (768, 2816)
    assert request_buckets("x", None) == (768, 640)


@pytest.mark.parametrize("sw", [1, 2, 3, 4])
def test_wav_decode_roundtrip(tmp_path, sw):
    from echo_tts_b200.audio import decode_file
    n, ch, sr = 1000, 2, 22_050
    x = (_sig(ch, n, seed=sw) * 0.3).clamp(-0.99, 0.99)
    full = float(1 << (8 * sw - 1))
    q = torch.round(x * full).to(torch.int64)
    frames = q.t().contiguous().view(-1).tolist()
    if sw == 1:
        raw = bytes(int(v) + 128 for v in frames)
    elif sw == 3:
        raw = b"".join(struct.pack("<i", int(v))[:3] for v in frames)
    else:
        raw = struct.pack("<%d%s" % (len(frames), {2: "h", 4: "i"}[sw]), *[int(v) for v in frames])
    p = str(tmp_path / f"t{sw}.wav")
    with wave.open(p, "wb") as w:
        w.setnchannels(ch); w.setsampwidth(sw); w.setframerate(sr); w.writeframes(raw)
    samples, rate = decode_file(p, max_duration=300)
    assert rate == sr and samples.shape == (ch, n)
    assert torch.equal(samples, q.to(torch.float32) / full)
    short, _ = decode_file(p, max_duration=0.01)
    assert short.shape == (ch, int(0.01 * sr))


@pytest.mark.gpu
@pytest.mark.parametrize("orig,new", RATES)
@pytest.mark.parametrize("channels,n", [(1, 7001), (2, 100_003), (3, 130)])
def test_gpu_prepare_audio_vs_oracle(orig, new, channels, n):
    from oracle import echo_oracle as O
    from echo_tts_b200.audio import prepare_audio
    x = 1.3 * _sig(channels, n, seed=n)
    ref = O.load_audio_post(x, orig)
    got = prepare_audio(x, orig).cpu()
    assert got.shape == ref.shape and got.dtype == torch.float32
    assert rel_l2(got, ref) < 2e-6                     # fp32 sums in another order
    assert float((got - ref).abs().max()) < 2e-6
    quiet = prepare_audio(0.1 * x, orig).cpu()         # peak < 1: never amplified (inference.py:147)
    assert rel_l2(quiet, O.load_audio_post(0.1 * x, orig)) < 2e-6


@pytest.mark.gpu
def test_gpu_load_audio_wav_and_pipeline_shapes(tmp_path):
    from oracle import echo_oracle as O
    from echo_tts_b200.inference import load_audio
    n, sr = 48_000, 48_000
    x = (_sig(2, n, seed=1) * 0.3).clamp(-0.99, 0.99)
    q = torch.round(x * 32768.0).to(torch.int16)
    p = str(tmp_path / "clip.wav")
    with wave.open(p, "wb") as w:
        w.setnchannels(2); w.setsampwidth(2); w.setframerate(sr); w.writeframes(q.t().contiguous().numpy().tobytes())
    got = load_audio(p)
    assert got.is_cuda and got.shape == (1, 44_100)
    assert rel_l2(got.cpu(), O.load_audio_post(q.float() / 32768.0, sr)) < 2e-6
