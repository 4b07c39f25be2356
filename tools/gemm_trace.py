#!/usr/bin/env python
"""In-kernel timeline of the CTA-pair GEMM (clock64 stamps, see echo_debug_gemm_trace in include/echo_b200.h).

    python tools/gemm_trace.py [--m 1920] [--case wo]
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

BF16 = torch.bfloat16
NAMES = ["setup", "first operands", "tile-0 MMAs issued", "tile-0 accumulator ready (epilogue view)",
         "tile-0 epilogue", "all roles done", "teardown"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--m", type=int, default=1920)
    ap.add_argument("--case", default="wo,w2,qkvg,w13")
    ap.add_argument("--seq", type=int, default=640)
    args = ap.parse_args()
    from echo_tts_b200 import ops, _lib as L
    from echo_tts_b200.model import rope_table
    lib = L.load()
    D, F_, H, S = 2048, 5888, 16, args.seq
    M = args.m
    g = torch.Generator(device="cuda").manual_seed(0)
    rnd = lambda *s, sc=1.0: (torch.randn(s, device="cuda", generator=g) * sc)
    xn, hm, x, gate = rnd(M, D).to(BF16), rnd(M, F_).to(BF16), rnd(M, D), rnd(M // S, D)
    qkvg = torch.empty((4, M, D), device="cuda", dtype=BF16)
    hmid = torch.empty((M, F_), device="cuda", dtype=BF16)
    qn = torch.ones(D, device="cuda")
    rope = rope_table(128, 4096, "cuda")
    W = dict(qkvg=rnd(4 * D, D, sc=D ** -0.5).to(BF16), wo=rnd(D, D, sc=D ** -0.5).to(BF16),
             w13=rnd(2 * F_, D, sc=D ** -0.5).to(BF16), w2=rnd(D, F_, sc=F_ ** -0.5).to(BF16))
    fns = dict(
        qkvg=lambda: ops.gemm_sections(xn, W["qkvg"], [(L.SEC_NORM_ROPE, qkvg[0], qn), (L.SEC_NORM_ROPE, qkvg[1], qn),
                                                       (L.SEC_PLAIN, qkvg[2], None), (L.SEC_SIGMOID, qkvg[3], None)],
                                       head_dim=128, rope_heads=H // 2, rope=rope, rows_per_seq=S, eps=1e-5),
        wo=lambda: ops.gemm_generic(xn, W["wo"], gate=gate, rows_per_gate=S, resid=x, out_f32=x),
        w13=lambda: ops.gemm_swiglu(xn, W["w13"], hmid),
        w2=lambda: ops.gemm_generic(hm, W["w2"], gate=gate, rows_per_gate=S, resid=x, out_f32=x),
        wo_split=lambda: ops.gemm_generic(xn, W["wo"], gate=gate, rows_per_gate=S, split_out=parts),
        w2_split=lambda: ops.gemm_generic(hm, W["w2"], gate=gate, rows_per_gate=S, split_out=parts))
    # S1-DAC ResidualUnit launches at the decoder's C = 192 stage (6 tiles per CTA pair): --case ae_k7,ae_k1,ae_k1_bf16
    Ta, Ca = 74 * 6 * 256, 192
    if "ae_" in args.case:
        s1a = rnd(Ta, Ca).to(BF16)
        w7a, w1a = rnd(Ca, 7 * Ca, sc=(7 * Ca) ** -0.5).to(BF16), rnd(Ca, Ca, sc=Ca ** -0.5).to(BF16)
        ba, ala, xa = rnd(Ca), rnd(Ca).abs() + 0.5, rnd(Ta, Ca)
        s2a, s3a = torch.empty((Ta, Ca), device="cuda", dtype=BF16), torch.empty((Ta, Ca), device="cuda", dtype=BF16)
        xoa = torch.empty((Ta, Ca), device="cuda")
        fns.update(
            ae_k7=lambda: ops.gemm_generic(s1a, w7a, taps=tuple(-(6 - k) * 9 for k in range(7)), bias=ba, out2_bf16=s2a, snake_alpha=ala),
            ae_k1=lambda: ops.gemm_generic(s2a, w1a, bias=ba, resid=xa, out_f32=xoa, out2_bf16=s3a, snake_alpha=ala),
            ae_k1_bf16=lambda: ops.gemm_generic(s2a, w1a, bias=ba, out_bf16=s3a))
    parts = torch.empty((3, M, D), device="cuda")
    buf = torch.zeros((512, 32), device="cuda", dtype=torch.int64)
    for name in args.case.split(","):
        fn = fns[name]
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        buf.zero_()
        os.environ["ECHO_GEMM_EPITRACE"] = "256"             # lean-epilogue chunk stamps in rows [256, 512)
        L.check(lib.echo_debug_gemm_trace(buf.data_ptr()))   # needs the debug build: make -C echo_tts_b200/csrc DEBUG=1
        fn()
        torch.cuda.synchronize()
        lib.echo_debug_gemm_trace(None)
        t = buf.cpu()
        et, t = t[256:], t[:256]
        nz = t[:, 0] != 0
        et = et[nz][0::2]
        t = t[nz]
        ld, fo = t[0::2], t[1::2]          # leader / follower CTA of each pair
        med = lambda v: float(v.float().median())
        print(f"== {name} M={M}: {t.shape[0]} CTAs; SM clocks, median over pairs")
        print(f"   setup {med(ld[:, 1] - ld[:, 0]):8.0f}   first operands after setup {med(ld[:, 2] - ld[:, 1]):8.0f}   "
              f"whole CTA {med(ld[:, 6] - ld[:, 0]):8.0f}   teardown {med(ld[:, 7] - ld[:, 6]):8.0f}")
        if bool((ld[:, 3] != 0).all()) and bool((ld[:, 4] != 0).all()) and os.environ.get("ECHO_STREAM_K", "0") != "1":
            print(f"   after setup: dependency resolved +{med(ld[:, 3] - ld[:, 1]):6.0f}  first stage requested +{med(ld[:, 4] - ld[:, 3]):6.0f}  "
                  f"landed (MMA thread's view) +{med(ld[:, 2] - ld[:, 4]):6.0f}")
        if name.startswith("qkvg") and bool((et[:, 0] != 0).any()):
            e = et[et[:, 0] != 0]
            for k, lab in ((0, "first tile"), (16, "second tile")):
                if bool((e[:, k + 8] != 0).all()):
                    d = lambda a, b: med(e[:, k + a] - e[:, k + b])
                    print(f"   sections epilogue, leader warp 2, {lab}: " + "  ".join(
                        f"head {h}: aux issued {d(1 + 4 * h, 4 * h):5.0f} ssq {d(2 + 4 * h, 1 + 4 * h):5.0f} "
                        f"chunk {d(3 + 4 * h, 2 + 4 * h):5.0f} chunk {d(4 + 4 * h, 3 + 4 * h):5.0f}" for h in (0, 1))
                        + f"  total {d(8, 0):6.0f}")
        elif bool((et[:, 0] != 0).any()):
            e = et[et[:, 0] != 0]
            d = lambda a, b: med(e[:, a] - e[:, b])
            row = "   lean epilogue, leader warp 2, per chunk (tmem ld | math | staged+fenced | bulk issued): "
            for i in range(4):
                if bool((e[:, 4 + 4 * i] != 0).all()):
                    prev = 0 if i == 0 else 4 * i
                    row += f"[{d(1 + 4 * i, prev):5.0f} {d(2 + 4 * i, 1 + 4 * i):5.0f} {d(3 + 4 * i, 2 + 4 * i):5.0f} {d(4 + 4 * i, 3 + 4 * i):5.0f}] "
            last = max(i for i in range(4) if bool((e[:, 4 + 4 * i] != 0).all()))
            print(row + f" drain {d(17, 4 + 4 * last):5.0f}  total {d(17, 0):6.0f}")
        mx = lambda v: float(v.float().max())
        sk = ld[ld[:, 3] != 0]
        if sk.shape[0] and os.environ.get("ECHO_STREAM_K", "0") == "1":
            last = torch.stack([sk[i, 8 + 4 * tc + 2: 8 + 4 * tc + 4] for i in range(sk.shape[0])
                                for tc in [max(k for k in range(6) if sk[i, 9 + 4 * k] != 0)]])
            print(f"   stream-K head units ({sk.shape[0]} pairs): sideload start @{med(sk[:, 3] - sk[:, 0]):8.0f}  took "
                  f"{med(sk[:, 4] - sk[:, 3]):8.0f} (max {mx(sk[:, 4] - sk[:, 3]):8.0f})  own MMAs done @{med(last[:, 0] - sk[:, 0]):8.0f}  "
                  f"merge+barrier {med(sk[:, 5] - last[:, 0]):8.0f}  epilogue {med(last[:, 1] - sk[:, 5]):8.0f}  end @{med(last[:, 1] - sk[:, 0]):8.0f} "
                  f"(max {mx(last[:, 1] - sk[:, 0]):8.0f})")
        for tc in range(6):
            b = 8 + 4 * tc
            ok = ld[:, b + 1] != 0
            if not bool(ok.any()):
                break
            l, f = ld[ok], fo[ok]
            print(f"   tile {tc} ({int(ok.sum())} pairs): acc free @{med(l[:, b] - l[:, 0]):8.0f}  MMAs issued @{med(l[:, b + 1] - l[:, 0]):8.0f} "
                  f"(issue span {med(l[:, b + 1] - l[:, b]):8.0f})  epilogue start @{med(l[:, b + 2] - l[:, 0]):8.0f}  "
                  f"duration leader {med(l[:, b + 3] - l[:, b + 2]):8.0f} follower {med(f[:, b + 3] - f[:, b + 2]):8.0f}")


if __name__ == "__main__":
    main()
