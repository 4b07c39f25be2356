#!/usr/bin/env python
"""One launch each of the S1-DAC k = 1 residual-unit convolution (C = 192) and its k = 7 neighbour, for ncu."""
import sys
sys.path.insert(0, "/root/repo")
import torch
from echo_tts_b200 import ops
BF16 = torch.bfloat16
g = torch.Generator(device="cuda").manual_seed(0)
rnd = lambda *s, sc=1.0: torch.randn(s, device="cuda", generator=g) * sc
T, C = 74 * 12 * 256, 192
s1, s2, s3 = rnd(T, C).to(BF16), rnd(T, C).to(BF16), torch.empty((T, C), device="cuda", dtype=BF16)
w7, w1 = rnd(C, 7 * C, sc=(7 * C) ** -0.5).to(BF16), rnd(C, C, sc=C ** -0.5).to(BF16)
b, alpha, x, xo = rnd(C), rnd(C).abs() + 0.5, rnd(T, C), torch.empty((T, C), device="cuda")
for _ in range(2):
    ops.gemm_generic(s2, w1, bias=b, resid=x, out_f32=xo, out2_bf16=s3, snake_alpha=alpha)
    ops.gemm_generic(s1, w7, taps=tuple(-(6 - k) * 9 for k in range(7)), bias=b, out2_bf16=s2, snake_alpha=alpha)
torch.cuda.synchronize()
print("ok")
