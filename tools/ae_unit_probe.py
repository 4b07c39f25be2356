#!/usr/bin/env python
"""S1-DAC ResidualUnit GEMMs alone (autoencoder.py:879-900) at the decoder's two last stages: the dilated k=7 conv
(Snake -> bf16) and the k=1 conv (bias + fp32 residual -> fp32 stream + Snake bf16), with parts of the k=1 epilogue
left out to see what bounds it.  Events around graph replays; buffers are > L2 so every launch streams from HBM."""
import sys
sys.path.insert(0, "/root/repo")
import torch
from echo_tts_b200 import ops

BF16, F32 = torch.bfloat16, torch.float32
g = torch.Generator(device="cuda").manual_seed(0)
rnd = lambda *s, sc=1.0: torch.randn(s, device="cuda", generator=g) * sc


def timed(fn, n=10):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr):
        for _ in range(n):
            fn()
    gr.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); gr.replay(); e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / n


for T, C, dil in ((655360, 192, 9), (1310720, 96, 9)):
    s1 = rnd(T, C).to(BF16)
    w7 = rnd(C, 7 * C, sc=(7 * C) ** -0.5).to(BF16)
    w1 = rnd(C, C, sc=C ** -0.5).to(BF16)
    b = rnd(C)
    alpha = rnd(C).abs() + 0.5
    x = rnd(T, C)
    s2 = torch.empty((T, C), device="cuda", dtype=BF16)
    xo = torch.empty((T, C), device="cuda")
    s1o = torch.empty((T, C), device="cuda", dtype=BF16)
    taps = tuple(-(6 - k) * dil for k in range(7))
    cases = {
        "k7 conv -> snake bf16": lambda: ops.gemm_generic(s1, w7, taps=taps, bias=b, out2_bf16=s2, snake_alpha=alpha),
        "k1 full (resid, f32 out, snake bf16)": lambda: ops.gemm_generic(s2, w1, bias=b, resid=x, out_f32=xo, out2_bf16=s1o, snake_alpha=alpha),
        "k1 in place (resid = out)": lambda: ops.gemm_generic(s2, w1, bias=b, resid=x, out_f32=x, out2_bf16=s1o, snake_alpha=alpha),
        "k1 no resid": lambda: ops.gemm_generic(s2, w1, bias=b, out_f32=xo, out2_bf16=s1o, snake_alpha=alpha),
        "k1 no snake out": lambda: ops.gemm_generic(s2, w1, bias=b, resid=x, out_f32=xo),
        "k1 snake out only": lambda: ops.gemm_generic(s2, w1, bias=b, out2_bf16=s1o, snake_alpha=alpha),
        "k1 bf16 out only": lambda: ops.gemm_generic(s2, w1, bias=b, out_bf16=s1o),
    }
    byt = {"k7 conv -> snake bf16": 4, "k1 full (resid, f32 out, snake bf16)": 12, "k1 in place (resid = out)": 12, "k1 no resid": 8,
           "k1 no snake out": 10, "k1 snake out only": 4, "k1 bf16 out only": 4}
    for name, fn in cases.items():
        us = timed(fn)
        print(f"T={T} C={C:3d} {name:40s} {us:8.1f} us  {T * C * byt[name] / us / 1e3:7.0f} GB/s algorithmic", flush=True)
