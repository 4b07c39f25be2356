#!/usr/bin/env python
"""Micro-benchmark of the DiT-layer GEMMs (through the C ABI) at the CFG-step (M=1920) and plain-step (M=640)
shapes, and of the attention kernel.  CUDA events on the launch stream; each timed loop cycles through `nw`
distinct weight sets (> L2) so weights stream from HBM as in the real step.

    python tools/gemm_bench.py [--m 1920 640] [--reps 5] [--only w13] [--quick]
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

BF16, F32 = torch.bfloat16, torch.float32


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--m", type=int, nargs="+", default=[1920, 640])
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--nw", type=int, default=8)
    ap.add_argument("--only", default="")
    ap.add_argument("--seq", type=int, default=640, help="latents per row (160 = a streaming block)")
    ap.add_argument("--bn", type=int, default=0, help="force block_n of the four projections (0 = library's choice)")
    ap.add_argument("--quick", action="store_true", help="1 rep, 1 weight set (for ncu)")
    ap.add_argument("--splits-wo", type=int, default=0, help="split-K units per tile of wo (0 = the model's choice)")
    ap.add_argument("--splits-w2", type=int, default=0, help="split-K units per tile of w2 (0 = the model's choice)")
    ap.add_argument("--drop", default="", help="comma list of launches left out of `seq` (norm1,qkvg,attn,wo,norm2,w13,w2): the "
                    "marginal cost of a kernel inside the layer sequence")
    args = ap.parse_args()
    from echo_tts_b200 import ops, _lib as L
    D, F_, H = 2048, 5888, 16
    nw = 1 if args.quick else args.nw
    reps = 1 if args.quick else args.reps
    g = torch.Generator(device="cuda").manual_seed(0)
    rnd = lambda *s, sc=1.0: (torch.randn(s, device="cuda", generator=g) * sc)
    W = dict(qkvg=[rnd(4 * D, D, sc=D ** -0.5).to(BF16) for _ in range(nw)],
             wo=[rnd(D, D, sc=D ** -0.5).to(BF16) for _ in range(nw)],
             w13=[rnd(2 * F_, D, sc=D ** -0.5).to(BF16) for _ in range(nw)],
             w2=[rnd(D, F_, sc=F_ ** -0.5).to(BF16) for _ in range(nw)])
    qn, kn = torch.ones(D, device="cuda"), torch.ones(D, device="cuda")
    from echo_tts_b200.model import rope_table
    rope = rope_table(128, 4096, "cuda")
    for M in args.m:
        S = args.seq
        R = M // S
        xn = rnd(M, D).to(BF16)
        hm = rnd(M, F_).to(BF16)
        x = rnd(M, D)
        gate = rnd(R, D)
        qkvg = torch.empty((4, M, D), device="cuda", dtype=BF16)
        ao = torch.empty((M, D), device="cuda", dtype=BF16)
        hmid = torch.empty((M, F_), device="cuda", dtype=BF16)
        kt, vt = rnd(1, 768, H, 128).to(BF16), rnd(1, 768, H, 128).to(BF16)
        ks, vs = rnd(1, 53, H, 128).to(BF16), rnd(1, 53, H, 128).to(BF16)
        tm = torch.zeros((R, 768), device="cuda", dtype=torch.uint8); tm[:, :256] = 1
        sm = torch.ones((R, 53), device="cuda", dtype=torch.uint8)
        if R == 3:
            tm[1] = 0; sm[2] = 0
        te, se = ops.mask_extent(tm), ops.mask_extent(sm)

        def f_qkvg(i):
            ops.gemm_sections(xn, W["qkvg"][i % nw], [(L.SEC_NORM_ROPE, qkvg[0], qn), (L.SEC_NORM_ROPE, qkvg[1], kn),
                                                      (L.SEC_PLAIN, qkvg[2], None), (L.SEC_SIGMOID, qkvg[3], None)],
                              head_dim=128, rope_heads=H // 2, rope=rope, rows_per_seq=S, eps=1e-5, block_n=args.bn)

        def f_wo(i):
            ops.gemm_generic(ao, W["wo"][i % nw], gate=gate, rows_per_gate=S, resid=x, out_f32=x, block_n=args.bn)

        ns = max(1, min(8 if M <= 256 else 4, 74 // (-(-M // 256) * 8)))       # model.EchoDiT.k_splits
        ns_wo, ns_w2 = args.splits_wo or ns, args.splits_w2 or ns
        if max(ns_wo, ns_w2) > 1:
            ns = max(ns, 2)
        parts = torch.empty((max(ns_wo, ns_w2, ns), M, D), device="cuda")

        def f_wo_split(i):
            ops.gemm_generic(ao, W["wo"][i % nw], gate=gate, rows_per_gate=S, split_out=parts[:ns_wo])

        def f_w2_split(i):
            ops.gemm_generic(hm, W["w2"][i % nw], gate=gate, rows_per_gate=S, split_out=parts[:ns_w2])

        def f_addnorm(i):
            ops.rmsnorm_mod(x, gate[0], gate[0], xn, 1e-5, S, gate.stride(0), delta=parts[:ns_wo])

        def f_addnorm2(i):
            ops.rmsnorm_mod(x, gate[0], gate[0], xn, 1e-5, S, gate.stride(0), delta=parts[:ns_w2])

        def f_w13(i):
            ops.gemm_swiglu(xn, W["w13"][i % nw], hmid, block_n=args.bn)

        def f_w2(i):
            ops.gemm_generic(hm, W["w2"][i % nw], gate=gate, rows_per_gate=S, resid=x, out_f32=x, block_n=args.bn)

        def f_attn(i):
            segs = [ops.Segment(qkvg[1].view(R, S, H, 128), qkvg[2].view(R, S, H, 128)),
                    ops.Segment(kt, vt, kv_batch_mod=1, mask=tm, extent=te),
                    ops.Segment(ks, vs, kv_batch_mod=1, mask=sm, extent=se)]
            ops.attention(qkvg[0].view(R, S, H, 128), segs, ao, gate=qkvg[3])

        def f_norm(i):
            ops.rmsnorm_mod(x, gate[0], gate[0], xn, 1e-5, S, gate.stride(0))

        drop = set(args.drop.split(",")) if args.drop else set()

        def f_seq(i):      # one decoder layer in launch order: cross-kernel effects (instruction cache, L2 state, PDL hand-offs)
            if "norm1" not in drop:
                f_addnorm2(i) if ns > 1 else f_norm(i)
            if "qkvg" not in drop:
                f_qkvg(i)
            if "attn" not in drop:
                f_attn(i)
            if "wo" not in drop:
                f_wo_split(i) if ns > 1 else f_wo(i)
            if "norm2" not in drop:
                f_addnorm(i) if ns > 1 else f_norm(i)
            if "w13" not in drop:
                f_w13(i)
            if "w2" not in drop:
                f_w2_split(i) if ns > 1 else f_w2(i)

        def f_seq2(i):     # the same launches, each kernel twice in a row (2 layers' worth): isolates per-kernel cold-start cost
            for f in ([f_norm, f_qkvg, f_attn] + ([f_wo_split, f_addnorm] if ns > 1 else [f_wo, f_norm]) + [f_w13] +
                      ([f_w2_split] if ns > 1 else [f_w2])):
                f(2 * i); f(2 * i + 1)

        keys = S + 256 + 53
        valid = [keys, S + 53, S + 256][:R] if R == 3 else [keys] * R
        cases = [("qkvg", f_qkvg, 2.0 * M * 4 * D * D), ("wo", f_wo, 2.0 * M * D * D), ("w13", f_w13, 2.0 * M * 2 * F_ * D),
                 ("w2", f_w2, 2.0 * M * D * F_), ("attn", f_attn, sum(4.0 * S * lv * D for lv in valid)),
                 ("norm", f_norm, 0.0)]
        cases += [("seq", f_seq, 0.0), ("seq2", f_seq2, 0.0)]
        if ns > 1:
            cases += [(f"wo/s{ns_wo}", f_wo_split, 2.0 * M * D * D), (f"w2/s{ns_w2}", f_w2_split, 2.0 * M * D * F_),
                      ("addnorm", f_addnorm, 0.0)]
        tot_us, tot_fl = 0.0, 0.0
        for name, fn, fl in cases:
            if args.only and name not in args.only.split(","):
                continue
            if not args.quick:
                for i in range(3):
                    fn(i)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            n = reps * nw
            if args.quick:
                e0.record()
                for i in range(n):
                    fn(i)
                e1.record()
            else:   # replay the launches from a CUDA graph: no Python/ctypes launch overhead between kernels
                gr = torch.cuda.CUDAGraph()
                with torch.cuda.graph(gr):
                    for i in range(n):
                        fn(i)
                gr.replay()
                torch.cuda.synchronize()
                e0.record()
                gr.replay()
                e1.record()
            torch.cuda.synchronize()
            us = e0.elapsed_time(e1) * 1e3 / n
            if name not in ("seq", "seq2"):
                tot_us += us; tot_fl += fl
            extra = f"{fl / us / 1e6:8.1f} TFLOP/s" if fl else f"{M * D * 6 / us / 1e3:8.1f} GB/s"
            print(f"M={M:5d} {name:5s} {us:8.2f} us  {extra}", flush=True)
        if tot_us > 0:
            print(f"M={M:5d} layer {tot_us:8.2f} us  {tot_fl / tot_us / 1e6:8.1f} TFLOP/s  (x24 = {tot_us * 24 / 1e3:.2f} ms/step)", flush=True)


if __name__ == "__main__":
    main()
