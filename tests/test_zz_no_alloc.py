"""Allocation-free split-K GEMM (``echo_gemm_bf16_splitk_noalloc``, csrc/noalloc.cu; no counterpart in the reference):
called through the C ABI, checked against the existing entry point
(``echo_gemm_bf16`` without k_splits) on two calls in a row -- the static scratch keeps the first call's contents, the second
call must not see them.  Collected last on purpose."""
import ctypes as C

import pytest
import torch

from conftest import rel_l2

pytestmark = pytest.mark.gpu
BF16, F32 = torch.bfloat16, torch.float32


def _desc(L, a, w, *, bias, gate, rows_per_gate, resid, out, k_splits=0):
    """echo_gemm_desc over preallocated tensors (no workspace lent: nothing here allocates)."""
    d = L.GemmDesc()
    d.a, d.b = a.data_ptr(), w.data_ptr()
    d.lda, d.ldb = a.stride(0), w.stride(0)
    d.rows, d.batch, d.a_batch_stride = a.shape[0], 1, a.shape[0] * a.stride(0)
    d.n, d.k_tap, d.taps = w.shape[0], a.shape[1], 1
    d.epilogue, d.act, d.impl = L.EPI_GENERIC, L.ACT_NONE, 2
    d.bias = bias.data_ptr()
    d.gate, d.gate_ld, d.rows_per_gate = gate.data_ptr(), gate.stride(0), rows_per_gate
    if resid is not None:
        d.resid, d.ld_resid = resid.data_ptr(), resid.stride(0)
    d.out_f32, d.ld_out_f32 = out.data_ptr(), out.stride(0)
    d.k_splits = k_splits
    return d


def test_splitk_noalloc_equals_plain_gemm_twice():
    from echo_tts_b200 import _lib as L
    lib = L.load()
    assert lib.echo_noalloc_scratch_bytes() == 32 << 20
    g = torch.Generator(device="cuda").manual_seed(5)
    rnd = lambda *s, sc=1.0: torch.randn(s, device="cuda", generator=g) * sc
    # call 1: the un-guided DiT step's w2 projection (640 tokens, K = 5888, 3 splits = 15.7 MB of partial products);
    # call 2: a smaller, ragged problem with other data, so stale partials of call 1 lie beyond and under its own
    for M, N, K, S, seq in ((640, 2048, 5888, 3, 640), (200, 384, 320, 2, 100)):
        a, w = rnd(M, K).to(BF16), rnd(N, K, sc=K ** -0.5).to(BF16)
        bias, gate, x0 = rnd(N), rnd(M // seq, N), rnd(M, N)
        ref, out = x0.clone(), x0.clone()
        d_ref = _desc(L, a, w, bias=bias, gate=gate, rows_per_gate=seq, resid=ref, out=ref)
        L.check(lib.echo_gemm_bf16(C.byref(d_ref), L.stream()))              # existing entry point: x += gate * (a w^T + b)
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated()
        d_new = _desc(L, a, w, bias=bias, gate=gate, rows_per_gate=seq, resid=out, out=out, k_splits=S)
        L.check(lib.echo_gemm_bf16_splitk_noalloc(C.byref(d_new), L.stream()))
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == before                       # nothing allocated through torch either
        want = x0 + gate.repeat_interleave(seq, 0) * (a.float() @ w.float().t() + bias)
        assert rel_l2(out, ref) < 2e-5, (M, N, K, rel_l2(out, ref))          # the tolerance of test_gemm_plain
        assert rel_l2(out, want) < 2e-5
        # without a residual the output is written, not accumulated
        out2 = torch.full((M, N), float("nan"), device="cuda")
        d2 = _desc(L, a, w, bias=bias, gate=gate, rows_per_gate=seq, resid=None, out=out2, k_splits=S)
        L.check(lib.echo_gemm_bf16_splitk_noalloc(C.byref(d2), L.stream()))
        assert rel_l2(out2, want - x0) < 2e-5


def test_splitk_noalloc_rejects_oversize_and_bad_arguments():
    from echo_tts_b200 import _lib as L
    lib = L.load()
    M, N, K = 1920, 2048, 256
    a = torch.zeros(M, K, device="cuda", dtype=BF16)
    w = torch.zeros(N, K, device="cuda", dtype=BF16)
    bias, gate, out = torch.zeros(N, device="cuda"), torch.ones(1, N, device="cuda"), torch.zeros(M, N, device="cuda")
    mk = lambda s: _desc(L, a, w, bias=bias, gate=gate, rows_per_gate=M, resid=None, out=out, k_splits=s)
    assert lib.echo_gemm_bf16_splitk_noalloc(C.byref(mk(3)), L.stream()) < 0      # 3 * 1920 * 2048 * 4 = 47 MB > 32 MiB
    msg = lib.echo_last_error().decode()
    assert "exceed the static scratch" in msg and "33554432" in msg
    assert lib.echo_gemm_bf16_splitk_noalloc(C.byref(mk(1)), L.stream()) < 0      # not a split-K request
    assert "k_splits must be >= 2" in lib.echo_last_error().decode()
    assert lib.echo_gemm_bf16_splitk_noalloc(None, L.stream()) < 0
    assert lib.echo_gemm_bf16_splitk_noalloc(C.byref(mk(2)), L.stream()) == 0     # 31.5 MB: fits
    torch.cuda.synchronize()
