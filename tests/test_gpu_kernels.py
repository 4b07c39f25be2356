"""Per-kernel parity on a B200: each hand-written kernel (called through the C ABI) against a plain torch
fp32 reference of the same op on the same inputs.  Tolerances: bf16-operand GEMM/attention outputs are
compared with references computed from the SAME bf16-rounded operands in fp32, so only accumulation order and
the final bf16 rounding differ (rel-L2 <= 4e-3); fp32 kernels <= 1e-5."""
import math

import pytest
import torch

from conftest import rel_l2

pytestmark = pytest.mark.gpu
BF16, F32 = torch.bfloat16, torch.float32


@pytest.fixture(scope="module")
def ops():
    from echo_tts_b200 import ops as o
    return o


@pytest.fixture(scope="module")
def L():
    from echo_tts_b200 import _lib
    return _lib


def _rand(shape, seed, scale=1.0, dtype=F32):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return (torch.randn(shape, device="cuda", generator=g) * scale).to(dtype)


@pytest.mark.parametrize("impl", [2])
@pytest.mark.parametrize("M,N,K,bn", [(128, 256, 64, 0), (256, 512, 512, 256), (200, 384, 320, 128), (1920, 2048, 2048, 0),
                                      (64, 96, 96, 0), (640, 2048, 5888, 128), (130, 64, 128, 64), (700, 192, 192, 0),
                                      (5000, 8192, 256, 0), (1920, 11776, 2048, 256)])
def test_gemm_plain(ops, impl, M, N, K, bn):
    if impl == 1 and M * N * K > 2e10:
        pytest.skip("legacy kernel: large case not needed")
    a, w = _rand((M, K), 1, dtype=BF16), _rand((N, K), 2, K ** -0.5, BF16)
    bias = _rand((N,), 3)
    out = torch.full((M, N), float("nan"), device="cuda", dtype=F32)
    outb = torch.empty((M, N), device="cuda", dtype=BF16)
    ops.gemm_generic(a, w, bias=bias, out_f32=out, out_bf16=outb, impl=impl, block_n=bn)
    ref = a.float() @ w.float().t() + bias
    assert rel_l2(out, ref) < 2e-5, (impl, M, N, K)
    assert rel_l2(outb.float(), ref) < 4e-3


@pytest.mark.parametrize("impl", [2])
def test_gemm_generic_epilogue_full(ops, L, impl):
    M, N, K, S = 384, 256, 192, 128
    a, w = _rand((M, K), 4, dtype=BF16), _rand((N, K), 5, K ** -0.5, BF16)
    bias, gate, resid, alpha = _rand((N,), 6), _rand((M // S, N), 7), _rand((M, N), 8), _rand((N,), 9).abs() + 0.5
    x = resid.clone()
    sn = torch.empty((M, N), device="cuda", dtype=BF16)
    ops.gemm_generic(a, w, bias=bias, act=L.ACT_GELU, gate=gate, rows_per_gate=S, resid=x, out_f32=x, out2_bf16=sn,
                     snake_alpha=alpha, impl=impl)
    ref = torch.nn.functional.gelu(a.float() @ w.float().t() + bias) * gate.repeat_interleave(S, 0) + resid
    assert rel_l2(x, ref) < 2e-5
    assert rel_l2(sn.float(), ref + torch.sin(alpha * ref) ** 2 / (alpha + 1e-9)) < 4e-3


@pytest.mark.parametrize("impl", [2])
@pytest.mark.parametrize("dil,taps_n,C,N,B,T", [(1, 7, 128, 128, 2, 300), (9, 7, 96, 96, 1, 700), (3, 7, 64, 256, 3, 130)])
def test_gemm_causal_conv_taps(ops, impl, dil, taps_n, C, N, B, T):
    """implicit-GEMM causal dilated conv1d == F.conv1d on left-padded input (autoencoder.py:285-289)."""
    x = _rand((B, T, C), 10, dtype=BF16)
    w = _rand((N, C, taps_n), 11, (C * taps_n) ** -0.5, BF16)
    wp = w.permute(0, 2, 1).reshape(N, taps_n * C).contiguous()
    out = torch.empty((B * T, N), device="cuda", dtype=F32)
    ops.gemm_generic(x.view(B * T, C), wp, taps=tuple(-(taps_n - 1 - k) * dil for k in range(taps_n)), batch=B,
                     out_f32=out, impl=impl)
    xin = torch.nn.functional.pad(x.float().transpose(1, 2), ((taps_n - 1) * dil, 0))
    ref = torch.nn.functional.conv1d(xin, w.float(), dilation=dil).transpose(1, 2).reshape(B * T, N)
    assert rel_l2(out, ref) < 2e-5


@pytest.mark.parametrize("impl", [2])
def test_gemm_conv_transpose(ops, impl):
    """2-tap GEMM == causal ConvTranspose1d(k=2s, stride s) with right trim (autoencoder.py:310-316)."""
    from echo_tts_b200.autoencoder import _pack_convT
    B, T, Cin, Cout, s = 2, 100, 128, 64, 4
    x = _rand((B, T, Cin), 12, dtype=BF16)
    w = _rand((Cin, Cout, 2 * s), 13, (2 * Cin) ** -0.5, BF16)
    out = torch.empty((B * T, s * Cout), device="cuda", dtype=F32)
    ops.gemm_generic(x.view(B * T, Cin), _pack_convT(w.float(), s), taps=(0, -1), batch=B, out_f32=out, impl=impl)
    ref = torch.nn.functional.conv_transpose1d(x.float().transpose(1, 2), w.float(), stride=s)[..., : T * s]
    assert rel_l2(out.view(B, T * s, Cout), ref.transpose(1, 2)) < 2e-5


@pytest.mark.parametrize("impl", [2])
def test_gemm_swiglu(ops, impl):
    M, K, F_ = 256, 256, 384
    a = _rand((M, K), 14, dtype=BF16)
    w1, w3 = _rand((F_, K), 15, K ** -0.5, BF16), _rand((F_, K), 16, K ** -0.5, BF16)
    out = torch.empty((M, F_), device="cuda", dtype=BF16)
    ops.gemm_swiglu(a, ops.pack_w13(w1, w3), out, impl=impl)
    ref = torch.nn.functional.silu(a.float() @ w1.float().t()) * (a.float() @ w3.float().t())
    assert rel_l2(out.float(), ref) < 4e-3


@pytest.mark.parametrize("impl", [2])
@pytest.mark.parametrize("hd,H,rope_heads", [(128, 4, 2), (64, 4, 4)])
def test_gemm_sections_norm_rope(ops, L, impl, hd, H, rope_heads):
    from oracle import echo_oracle as O
    R, S, K, start = 2, 96, 256, 5
    D = H * hd
    M = R * S
    a = _rand((M, K), 17, dtype=BF16)
    w = _rand((4 * D, K), 18, K ** -0.5, BF16)
    qn, kn = _rand((D,), 19).abs() + 0.5, _rand((D,), 20).abs() + 0.5
    cos, sin = O.rope_table(hd, 4096)
    rope = (cos.cuda().contiguous(), sin.cuda().contiguous())
    outs = torch.empty((4, M, D), device="cuda", dtype=BF16)
    ops.gemm_sections(a, w, [(L.SEC_NORM_ROPE, outs[0], qn), (L.SEC_NORM_ROPE, outs[1], kn), (L.SEC_PLAIN, outs[2], None),
                             (L.SEC_SIGMOID, outs[3], None)], head_dim=hd, rope_heads=rope_heads, rope=rope,
                      pos_offset=start, rows_per_seq=S, eps=1e-5, impl=impl)
    acc = (a.float() @ w.float().t()).cpu()
    q, k, v, g = acc.split(D, dim=1)

    def nr(t, wn):
        t = O.rms_norm(t.reshape(R, S, H, hd), wn.cpu().reshape(H, hd), 1e-5)
        rot = O.rope_apply(t[:, :, :rope_heads], cos[start:start + S], sin[start:start + S])
        return torch.cat([rot, t[:, :, rope_heads:]], 2).reshape(M, D)
    for got, ref in zip(outs, (nr(q, qn), nr(k, kn), v, torch.sigmoid(g))):
        assert rel_l2(got.float().cpu(), ref) < 4e-3


@pytest.mark.parametrize("M,N,K", [(1920, 2048, 2048), (1920, 2048, 5888), (640, 2048, 2048), (640, 8192, 2048), (2304, 768, 1024)])
def test_gemm_stream_k_generic(ops, M, N, K):
    """Ragged tile counts are scheduled stream-K (partial accumulators exchanged through the lent workspace): same
    result as whole-tile scheduling up to fp32 summation order, identical from run to run (self-resetting flags)."""
    a, w = _rand((M, K), 41, dtype=BF16), _rand((N, K), 42, K ** -0.5, BF16)
    bias, gate, resid = _rand((N,), 43), _rand((3, N), 44), _rand((M, N), 45)
    rpg = (M + 2) // 3
    ref = (a.float() @ w.float().t() + bias) * gate[torch.arange(M, device="cuda") // rpg] + resid
    outs = []
    for impl in (3, 3, 2):                                   # impl 3: the stream-K schedule is allowed
        x = resid.clone()
        ops.gemm_generic(a, w, bias=bias, gate=gate, rows_per_gate=rpg, resid=x, out_f32=x, impl=impl)
        outs.append(x)
    assert torch.equal(outs[0], outs[1])
    assert rel_l2(outs[0], ref) < 2e-5 and rel_l2(outs[2], ref) < 2e-5
    assert rel_l2(outs[0], outs[2]) < 5e-6


def test_gemm_stream_k_fused_epilogues(ops, L):
    """SECTIONS (per-head norm needs the whole fixed-up accumulator row) and SWIGLU epilogues on stream-K tiles."""
    from oracle import echo_oracle as O
    R, S, K, H, hd = 1, 640, 2048, 16, 128
    D, M = H * hd, R * S
    a = _rand((M, K), 46, dtype=BF16)
    w = _rand((4 * D, K), 47, K ** -0.5, BF16)
    qn, kn = _rand((D,), 48).abs() + 0.5, _rand((D,), 49).abs() + 0.5
    cos, sin = O.rope_table(hd, 4096)
    rope = (cos.cuda().contiguous(), sin.cuda().contiguous())
    res = {}
    for sk, impl in (("1", 3), ("0", 2)):
        outs = torch.empty((4, M, D), device="cuda", dtype=BF16)
        ops.gemm_sections(a, w, [(L.SEC_NORM_ROPE, outs[0], qn), (L.SEC_NORM_ROPE, outs[1], kn), (L.SEC_PLAIN, outs[2], None),
                                 (L.SEC_SIGMOID, outs[3], None)], head_dim=hd, rope_heads=8, rope=rope, pos_offset=0,
                          rows_per_seq=S, eps=1e-5, impl=impl)
        res[sk] = outs.float()
    assert rel_l2(res["1"], res["0"]) < 2e-3          # bf16 outputs: an fp32 reordering flips a few roundings
    acc = (a.float() @ w.float().t())
    assert rel_l2(res["1"][2], acc[:, 2 * D:3 * D]) < 4e-3 and rel_l2(res["1"][3], torch.sigmoid(acc[:, 3 * D:])) < 4e-3
    F_ = 5888
    w1, w3 = _rand((F_, K), 50, K ** -0.5, BF16), _rand((F_, K), 51, K ** -0.5, BF16)
    w13 = ops.pack_w13(w1, w3)
    a2 = _rand((896, K), 52, dtype=BF16)               # 4 x 46 = 184 tiles on 74 pairs: ragged
    ref = torch.nn.functional.silu(a2.float() @ w1.float().t()) * (a2.float() @ w3.float().t())
    for impl in (3, 2):
        out = torch.empty((896, F_), device="cuda", dtype=BF16)
        ops.gemm_swiglu(a2, w13, out, impl=impl)
        assert rel_l2(out.float(), ref) < 4e-3, impl


@pytest.mark.parametrize("M,N,K,S", [(640, 2048, 5888, 3), (640, 2048, 2048, 3), (100, 256, 256, 4), (300, 512, 1024, 2)])
def test_gemm_split_k_deferred_residual(ops, M, N, K, S):
    """split-K partial products (gated) + the fused add-norm that consumes them == the one-pass residual GEMM."""
    a, w = _rand((M, K), 40, dtype=BF16), _rand((N, K), 41, K ** -0.5, BF16)
    gate, x0 = _rand((1, N), 42), _rand((M, N), 43)
    sc1, sh = _rand((N,), 44) + 1.0, _rand((N,), 45)
    parts = torch.full((S, M, N), float("nan"), device="cuda", dtype=F32)
    ops.gemm_generic(a, w, gate=gate, split_out=parts)
    ref_delta = (a.float() @ w.float().t()) * gate
    assert rel_l2(parts.sum(0), ref_delta) < 2e-5
    x = x0.clone()
    y = torch.empty((M, N), device="cuda", dtype=BF16)
    ops.rmsnorm_mod(x, sc1, sh, y, 1e-5, delta=parts)
    xr = x0 + ref_delta
    assert rel_l2(x, xr) < 2e-5
    ref = xr * torch.rsqrt(xr.pow(2).mean(-1, keepdim=True) + 1e-5) * sc1 + sh
    assert rel_l2(y.float(), ref) < 3e-3


def test_sgemm_and_batched(ops, L):
    a, w, b, add = _rand((70, 100), 21), _rand((90, 100), 22), _rand((90,), 23), _rand((70, 90), 24)
    out = torch.empty((70, 90), device="cuda")
    ops.sgemm(a, w, bias=b, addend=add, out=out, pre_silu=True, post=L.POST_TANH, out_scale=0.5)
    ref = torch.tanh((torch.nn.functional.silu(a) @ w.t() + b) * 0.5 + add)
    assert rel_l2(out, ref) < 1e-5
    az, wz, bz = _rand((5, 33, 64), 25), _rand((5, 48, 64), 26), _rand((5, 48), 27)
    outz = torch.empty((5, 33, 48), device="cuda")
    ops.sgemm(az, wz, bias=bz, out=outz, post=L.POST_PLUS1)
    assert rel_l2(outz, torch.einsum("zmk,znk->zmn", az, wz) + bz[:, None] + 1) < 1e-5


def test_sgemm_scalar_and_vector_paths_and_b_static_flag(ops, L):
    """fp32 sgemm: K not a multiple of 4 / misaligned rows take the scalar K-step-16 form, everything else the 16-byte
    prefetching form -- same results; echo_gemm_desc.b_static only moves WHEN the first weight tiles are requested."""
    import ctypes as C
    for (m, n, k) in ((40, 200, 81), (40, 200, 84), (3, 64, 2048), (130, 70, 36)):
        a, w = _rand((m, k), 70), _rand((n, k), 71, k ** -0.5)
        out = torch.empty((m, n), device="cuda")
        ops.sgemm(a, w, out=out, pre_silu=True)
        assert rel_l2(out, torch.nn.functional.silu(a) @ w.t()) < 1e-5, (m, n, k)
    a, w = _rand((300, 512), 72, dtype=BF16), _rand((384, 512), 73, 512 ** -0.5, BF16)
    outs = []
    for flag in (1, 0):
        d = ops._base_gemm(a, w)
        d.epilogue, d.b_static = L.EPI_GENERIC, flag
        o = torch.empty((300, 384), device="cuda", dtype=F32)
        d.out_f32, d.ld_out_f32 = o.data_ptr(), o.stride(0)
        L.check(L.load().echo_gemm_bf16(C.byref(d), L.stream()))
        outs.append(o)
    assert torch.equal(outs[0], outs[1]) and rel_l2(outs[0], a.float() @ w.float().t()) < 2e-5


def test_rmsnorm_mod(ops):
    M, D, S = 96, 2048, 32
    x = _rand((M, D), 28, 3.0)
    tab = _rand((M // S, 2, D), 29)
    y = torch.empty((M, D), device="cuda", dtype=BF16)
    ops.rmsnorm_mod(x, tab[0, 1], tab[0, 0], y, 1e-5, rows_per_mod=S, mod_ld=tab.stride(0))
    ref = x * torch.rsqrt(x.pow(2).mean(-1, keepdim=True) + 1e-5) * tab[:, 1].repeat_interleave(S, 0) + \
        tab[:, 0].repeat_interleave(S, 0)
    assert rel_l2(y.float(), ref) < 3e-3


# ---- narrow tail: the ragged last round of a launch runs as half-width tiles (gemm2.cu launch_pair_t) ------------------
@pytest.mark.parametrize("M,N,K", [(1920, 8192, 320), (640, 8192, 256), (2000, 5120, 192), (700, 8192, 64)])
def test_gemm_narrow_tail_matches_plain_schedule(ops, M, N, K):
    """Shapes whose 256-wide tiles leave a ragged last round that fits one round of 128-wide tiles (e.g. 8 x 32 = 256
    tiles on 74 CTA pairs: 3 rounds + 34).  Same per-element sums as the plain schedule -> bit-identical results, for
    the generic, sections (per-head RMSNorm + RoPE + sigmoid) and SwiGLU epilogues; and both match the fp32 reference."""
    from oracle import echo_oracle as O
    from echo_tts_b200 import _lib as L
    a = _rand((M, K), 61, dtype=BF16)
    w = _rand((N, K), 62, K ** -0.5, BF16)
    bias = _rand((N,), 63)
    D = N // 4
    qn, kn = _rand((D,), 64).abs() + 0.5, _rand((D,), 65).abs() + 0.5
    cos, sin = O.rope_table(128, 4096)
    rope = (cos.cuda().contiguous(), sin.cuda().contiguous())
    res = {}
    for tail, impl in (("1", 2), ("0", 4)):   # impl 4 = the plain schedule, no half-width last round
        out = torch.full((M, N), float("nan"), device="cuda", dtype=F32)
        ops.gemm_generic(a, w, bias=bias, out_f32=out, impl=impl)
        sec = torch.full((4, M, D), float("nan"), device="cuda", dtype=BF16)
        ops.gemm_sections(a, w, [(L.SEC_NORM_ROPE, sec[0], qn), (L.SEC_NORM_ROPE, sec[1], kn), (L.SEC_PLAIN, sec[2], None),
                                 (L.SEC_SIGMOID, sec[3], None)], head_dim=128, rope_heads=D // 256, rope=rope,
                          rows_per_seq=M, eps=1e-5, impl=impl)
        sw = torch.full((M, N // 2), float("nan"), device="cuda", dtype=BF16)
        ops.gemm_swiglu(a, w, sw, impl=impl)
        res[tail] = (out, sec, sw)
    for x, y in zip(res["1"], res["0"]):
        assert torch.equal(x, y)
    assert rel_l2(res["1"][0], a.float() @ w.float().t() + bias) < 2e-5


def test_gemm_narrow_tail_with_batches_and_taps(ops):
    """Narrow tail on a batched implicit-GEMM conv (6 row tiles x 4 column tiles x 4 batch entries = 96 tiles -> 74 + 22
    -> 44 half-width tiles): tile -> (row tile, column tile, batch) decoding of the half tiles, bit-identical to the plain
    schedule and equal to the fp32 conv."""
    B, T, C, N, dil = 4, 1500, 128, 1024, 3
    x = _rand((B, T, C), 71, dtype=BF16)
    w = _rand((N, C, 7), 72, (C * 7) ** -0.5, BF16)
    wp = w.permute(0, 2, 1).reshape(N, 7 * C).contiguous()
    outs = {}
    for tail, impl in (("1", 2), ("0", 4)):   # impl 4 = the plain schedule, no half-width last round
        out = torch.full((B * T, N), float("nan"), device="cuda", dtype=F32)
        ops.gemm_generic(x.view(B * T, C), wp, taps=tuple(-(6 - k) * dil for k in range(7)), batch=B, out_f32=out, impl=impl)
        outs[tail] = out
    assert torch.equal(outs["1"], outs["0"])
    xin = torch.nn.functional.pad(x.float().transpose(1, 2), (6 * dil, 0))
    ref = torch.nn.functional.conv1d(xin, w.float(), dilation=dil).transpose(1, 2).reshape(B * T, N)
    assert rel_l2(outs["1"], ref) < 2e-5


# ---- row-per-thread TMA epilogue of the GENERIC GEMM (gemm_epilogue_t.cuh epi_t_row; the S1-DAC convolutions) ----------
@pytest.mark.parametrize("T,C,N,taps,batch", [(5000, 192, 192, 1, 1), (3001, 96, 96, 7, 1), (700, 384, 384, 1, 3),
                                             (1500, 128, 1024, 7, 2), (257, 64, 40, 1, 1), (33, 256, 256, 1, 1)])
@pytest.mark.parametrize("mode", ["resid_f32_snake", "inplace_snake", "snake_only", "f32_only", "bf16_only", "f32_bf16"])
def test_gemm_row_epilogue_bit_identical_to_transposing_epilogue(ops, T, C, N, taps, batch, mode):
    """impl 2 takes the row / TMA epilogue for bias [+ residual] -> fp32 and / or one bf16 output, impl 5 the transposing
    one: same per-element arithmetic -> identical bits, for ragged row / column ends (TMA clipping), batches, taps,
    narrow tiles, the in-place residual and every output combination; and both match the fp32 reference."""
    dil = 3
    x = _rand((batch, T, C), 91, dtype=BF16)
    w = _rand((N, C, taps), 92, (C * taps) ** -0.5, BF16)
    wp = w.permute(0, 2, 1).reshape(N, taps * C).contiguous()
    tp = tuple(-(taps - 1 - k) * dil for k in range(taps))
    bias, alpha = _rand((N,), 93), _rand((N,), 94).abs() + 0.5
    r0 = _rand((batch * T, N), 95)
    outs = {}
    for impl in (2, 5):
        kw = dict(taps=tp, batch=batch, bias=bias, impl=impl)
        res = r0.clone()
        f32 = torch.full((batch * T, N), float("nan"), device="cuda")
        b16 = torch.full((batch * T, N), float("nan"), device="cuda", dtype=BF16)
        a = x.view(batch * T, C)
        if mode == "resid_f32_snake":
            ops.gemm_generic(a, wp, resid=res, out_f32=f32, out2_bf16=b16, snake_alpha=alpha, **kw)
        elif mode == "inplace_snake":
            ops.gemm_generic(a, wp, resid=res, out_f32=res, out2_bf16=b16, snake_alpha=alpha, **kw); f32 = res
        elif mode == "snake_only":
            ops.gemm_generic(a, wp, out2_bf16=b16, snake_alpha=alpha, **kw)
        elif mode == "f32_only":
            ops.gemm_generic(a, wp, resid=res, out_f32=f32, **kw)
        elif mode == "bf16_only":
            ops.gemm_generic(a, wp, out_bf16=b16, **kw)
        else:
            ops.gemm_generic(a, wp, out_f32=f32, out_bf16=b16, **kw)
        outs[impl] = (f32, b16)
    has_f32 = mode not in ("snake_only", "bf16_only")
    has_b16 = mode != "f32_only"
    if has_f32:
        assert torch.equal(outs[2][0], outs[5][0])
    if has_b16:
        assert torch.equal(outs[2][1].view(torch.int16), outs[5][1].view(torch.int16))
    xin = torch.nn.functional.pad(x.float().transpose(1, 2), ((taps - 1) * dil, 0))
    ref = torch.nn.functional.conv1d(xin, w.float(), dilation=dil).transpose(1, 2).reshape(batch * T, N) + bias
    if mode in ("resid_f32_snake", "inplace_snake", "f32_only"):
        ref = ref + r0
    if has_f32:
        assert rel_l2(outs[2][0], ref) < 3e-5
    if has_b16:
        want = ref + torch.sin(alpha * ref) ** 2 / (alpha + 1e-9) if "snake" in mode else ref
        assert rel_l2(outs[2][1].float(), want) < 5e-3
