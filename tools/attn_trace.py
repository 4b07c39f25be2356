#!/usr/bin/env python
"""In-kernel timeline of the tcgen05 attention kernel (see echo_debug_attn_trace)."""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

BF16 = torch.bfloat16


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=1)
    ap.add_argument("--seq", type=int, default=640, help="queries per row (160 = a streaming block)")
    ap.add_argument("--prefix", type=int, default=0, help="keys of a latent prefix segment (480 behind three blocks)")
    ap.add_argument("--nomask", action="store_true", help="no masks / extents: the tile list needs no global loads")
    args = ap.parse_args()
    from echo_tts_b200 import ops, _lib as L
    lib = L.load()
    R, S, H = args.rows, args.seq, 16
    g = torch.Generator(device="cuda").manual_seed(0)
    rnd = lambda *s: torch.randn(s, device="cuda", generator=g).to(BF16)
    qkvg = rnd(4, R, S, H, 128)
    ao = torch.empty((R * S, H * 128), device="cuda", dtype=BF16)
    kt, vt, ks, vs = rnd(1, 768, H, 128), rnd(1, 768, H, 128), rnd(1, 53, H, 128), rnd(1, 53, H, 128)
    tm = torch.zeros((R, 768), device="cuda", dtype=torch.uint8); tm[:, :256] = 1
    sm = torch.ones((R, 53), device="cuda", dtype=torch.uint8)
    if R == 3:
        tm[1] = 0; sm[2] = 0
    te, se = ops.mask_extent(tm), ops.mask_extent(sm)
    segs = [ops.Segment(qkvg[1], qkvg[2]), ops.Segment(kt, vt, kv_batch_mod=1, mask=tm, extent=te),
            ops.Segment(ks, vs, kv_batch_mod=1, mask=sm, extent=se)]
    if args.prefix:
        kp, vp = rnd(1, args.prefix, H, 128), rnd(1, args.prefix, H, 128)
        segs.insert(1, ops.Segment(kp, vp, kv_batch_mod=1))
    if args.nomask:
        segs = [ops.Segment(qkvg[1], qkvg[2]), ops.Segment(kt[:, :256], vt[:, :256], kv_batch_mod=1),
                ops.Segment(ks, vs, kv_batch_mod=1)]
    fn = lambda: ops.attention(qkvg[0], segs, ao, gate=qkvg[3].reshape(R * S, H * 128))
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    n = ((S + 127) // 128) * H * R * 4
    buf = torch.zeros((n, 64), device="cuda", dtype=torch.int64)
    L.check(lib.echo_debug_attn_trace(buf.data_ptr()))   # needs the debug build: make -C echo_tts_b200/csrc DEBUG=1
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    fn()
    e1.record()
    torch.cuda.synchronize()
    lib.echo_debug_attn_trace(None)
    t = buf.cpu()
    t = t[t[:, 0] != 0]
    n = t.shape[0]
    gs, ge = t[:, 57], t[:, 58]
    print(f"globaltimer: first CTA start -> last CTA exit {int(ge.max() - gs.min())} ns; CTA starts spread over "
          f"{int(gs.max() - gs.min())} ns; CTA lifetime median {int((ge - gs).median())} / p90 "
          f"{int((ge - gs).float().quantile(0.9))} / max {int((ge - gs).max())} ns; events around the launch "
          f"{e0.elapsed_time(e1) * 1e3:.1f} us")
    med = lambda v: float(v.float().median())
    t0 = t[:, 0]
    print(f"rows={R}: {n} CTAs, tiles/CTA median {med(t[:, 63]):.0f}; span first start -> last exit "
          f"{int(t[:, 62].max() - t0.min())} clk (globaltimer-free; SM clocks are per-SM, indicative only)")
    print(f"  warp 0 past PDL wait @{med(t[:, 2] - t0):7.0f}   tile list built @{med(t[:, 3] - t0):7.0f}")
    print(f"  extents loaded @{med(t[:, 56] - t0):7.0f}   MMA warp: in loop @{med(t[:, 52] - t0):7.0f}  Q landed @{med(t[:, 53] - t0):7.0f}  "
          f"K(0) landed @{med(t[:, 54] - t0):7.0f}  S(0) issued @{med(t[:, 55] - t0):7.0f}")
    print(f"  relative to CTA start, median over CTAs: setup done @{med(t[:, 1] - t0):7.0f}   last PV done "
          f"@{med(t[:, 60] - t0):7.0f}   epilogue done @{med(t[:, 61] - t0):7.0f}   exit @{med(t[:, 62] - t0):7.0f}")
    if bool((t[:, 48] != 0).any()) and med(t[:, 63]) < 11:
        e = t[t[:, 48] != 0]
        print(f"  split-KV (stamps in tile 11's slots), relative to last PV done: partial stored @{med(e[:, 48] - e[:, 60]):7.0f}  "
              f"fenced + CTA barrier @{med(e[:, 49] - e[:, 60]):7.0f}  all parts arrived @{med(e[:, 50] - e[:, 60]):7.0f}  "
              f"own columns merged + stored @{med(e[:, 51] - e[:, 60]):7.0f}")
    for j in range(12):
        b = 4 + 4 * j
        print(f"  tile {j:2d}: S ready @{med(t[:, b] - t0):7.0f}  PV(j-1) done @{med(t[:, b + 1] - t0):7.0f}  "
              f"P published @{med(t[:, b + 2] - t0):7.0f}  PV issued @{med(t[:, b + 3] - t0):7.0f}")


if __name__ == "__main__":
    main()
