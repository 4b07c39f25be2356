#!/usr/bin/env python
"""Profiling driver for the HBM-bound kernels added late in round 1: the S1-DAC output conv at full size
(1,310,720 samples x 96 channels) and the audio front-end (30 s of 48 kHz stereo -> 44.1 kHz mono + peak normalisation).

    ncu --set full --clock-control none -k regex:"conv_out_tanh|resample_mono|peak_normalize" python tools/profile_misc.py
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402


def main():
    from echo_tts_b200 import ops
    from echo_tts_b200.audio import prepare_audio
    g = torch.Generator(device="cuda").manual_seed(0)
    T, C = 640 * 2048, 96
    x = torch.randn((1, T, C), device="cuda", generator=g).to(torch.bfloat16)
    w = torch.randn((7, C), device="cuda", generator=g) * 0.05
    wav = torch.empty((1, T), device="cuda")
    pcm = torch.rand((2, 30 * 48000), device="cuda", generator=g) * 2 - 1
    for _ in range(2):
        ops.conv_out_tanh(x, w, 0.01, wav)
        prepare_audio(pcm, 48000)
    torch.cuda.synchronize()
    e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    e[0].record(); ops.conv_out_tanh(x, w, 0.01, wav); e[1].record(); prepare_audio(pcm, 48000); e[2].record()
    torch.cuda.synchronize()
    print(f"conv_out_tanh {e[0].elapsed_time(e[1]) * 1e3:.1f} us ({(T * C * 2 + T * 4) / e[0].elapsed_time(e[1]) / 1e6:.0f} GB/s), "
          f"prepare_audio {e[1].elapsed_time(e[2]) * 1e3:.1f} us")


if __name__ == "__main__":
    main()
