#!/usr/bin/env python
"""What does cuBLAS launch at the DiT projection shapes?  torch.matmul (bf16) at the four projections x M in {1920, 640};
run under `ncu --set full -k regex:.` to read kernel names (tile / cluster / stages), launch configuration and L2 / DRAM
traffic beside our own kernel's (profiles/r2_cublas_same_shape.md).

    python tools/cublas_probe.py [--m 1920 640]
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--m", type=int, nargs="+", default=[1920, 640])
    args = ap.parse_args()
    D, F = 2048, 5888
    g = torch.Generator(device="cuda").manual_seed(0)
    bf = torch.bfloat16
    shapes = dict(qkvg=(4 * D, D), wo=(D, D), w13=(2 * F, D), w2=(D, F))
    for M in args.m:
        for name, (N, K) in shapes.items():
            a = torch.randn(M, K, device="cuda", generator=g).to(bf)
            w = torch.randn(N, K, device="cuda", generator=g).to(bf)
            out = torch.empty(M, N, device="cuda", dtype=bf)
            for _ in range(2):
                torch.matmul(a, w.t(), out=out)
            torch.cuda.synchronize()
            torch.cuda.profiler.start()
            torch.matmul(a, w.t(), out=out)
            torch.cuda.synchronize()
            torch.cuda.profiler.stop()
            print(f"cublas {name} M={M} N={N} K={K} ok", flush=True)


if __name__ == "__main__":
    main()
