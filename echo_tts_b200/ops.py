"""Tensor-level wrappers over the C ABI (``include/echo_b200.h``).  Host plumbing only: every
function validates shapes, fills a descriptor and launches hand-written kernels on the current
CUDA stream.  Nothing here computes on the CPU."""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional, Sequence, Tuple

import torch

from . import _lib as L

BF16, F32 = torch.bfloat16, torch.float32
STREAM_K = os.environ.get("ECHO_STREAM_K") == "1"   # read once at import: allow the stream-K schedule (off by default)
GEMM_IMPL = 3 if STREAM_K else 2   # echo_gemm_desc.impl: the CTA-pair persistent tcgen05 kernel (the only GEMM kernel of
                                   # the library); 3 = the same with stream-K scheduling of ragged launches allowed


def _need(t: torch.Tensor, dtype, name: str):
    if not t.is_cuda:
        raise RuntimeError(f"{name}: CUDA tensor required (no CPU fallback)")
    if t.dtype != dtype:
        raise TypeError(f"{name}: expected {dtype}, got {t.dtype}")


_GEMM_WS = {}


def gemm_workspace(device: torch.device) -> torch.Tensor:
    """Stream-K scratch lent to every impl-2 GEMM on ``device`` (flags + one partial accumulator per CTA pair, zeroed
    once).  One buffer per device: GEMMs of this package run on one stream at a time per device."""
    key = device.index if device.index is not None else torch.cuda.current_device()
    ws = _GEMM_WS.get(key)
    if ws is None:
        ws = torch.zeros(int(L.load().echo_gemm_workspace_bytes()), device=device, dtype=torch.uint8)
        _GEMM_WS[key] = ws
    return ws


_ATTN_WS = {}
ATTN_WS_SLOT = 0     # which of the per-device exchange buffers attention launches use: work that runs on a second stream
                     # next to the main one (the speaker encoder, model.kv_caches_text_speaker) switches to slot 1


def attn_workspace(device: torch.device) -> torch.Tensor:
    """Split-KV exchange buffer lent to every tcgen05 attention launch on ``device`` (echo_attn_desc.workspace: arrival
    counters + 296 partial (O, m, l) = two CTAs per SM, zeroed once).  One buffer per (device, slot): launches that may
    overlap on two streams must not share one (``ATTN_WS_SLOT``).  Created at the first -- eager -- request of a shape,
    i.e. before any CUDA-graph capture needs it."""
    key = (device.index if device.index is not None else torch.cuda.current_device(), ATTN_WS_SLOT)
    ws = _ATTN_WS.get(key)
    if ws is None:
        ws = torch.zeros(int(L.load().echo_attention_workspace_bytes(1, 128 * 74, 1, 4)), device=device, dtype=torch.uint8)
        _ATTN_WS[key] = ws
    return ws


def _base_gemm(a: torch.Tensor, w: torch.Tensor, taps: Sequence[int] = (0,), batch: int = 1,
               impl: Optional[int] = None, block_n: int = 0, a_ctx: int = 0) -> L.GemmDesc:
    """a: bf16 [batch*rows, k_tap] (row-major, contiguous rows); w: bf16 [n, taps*k_tap].  ``a_ctx``: rows of left
    context stored in memory before a[0] (a view into a larger buffer; batch 1): read instead of zero padding."""
    _need(a, BF16, "gemm.a"); _need(w, BF16, "gemm.w")
    assert a.dim() == 2 and w.dim() == 2 and a.stride(1) == 1 and w.stride(1) == 1
    d = L.GemmDesc()
    total_rows, k_tap = a.shape
    assert total_rows % batch == 0
    d.a, d.b = a.data_ptr(), w.data_ptr()
    d.lda, d.ldb = a.stride(0), w.stride(0)
    d.rows, d.batch = total_rows // batch, batch
    d.a_batch_stride = d.rows * a.stride(0)
    d.n, d.k_tap, d.taps = w.shape[0], k_tap, len(taps)
    assert w.shape[1] == len(taps) * k_tap, (w.shape, taps, k_tap)
    for i, s in enumerate(taps):
        d.tap_shift[i] = s
    d.impl = GEMM_IMPL if impl is None else impl
    d.block_n = block_n
    assert a_ctx == 0 or batch == 1
    d.a_ctx = a_ctx
    d.rows_per_gate = 1
    d.b_static = 1          # every caller in this package passes load-time weights as `w` (include/echo_b200.h: b_static)
    if d.impl == 3:
        ws = gemm_workspace(a.device)
        d.workspace, d.workspace_bytes = ws.data_ptr(), ws.numel()
    return d


def gemm_generic(a, w, *, taps=(0,), batch=1, bias=None, act=L.ACT_NONE, gate=None, rows_per_gate=0,
                 resid=None, out_f32=None, out_bf16=None, out2_bf16=None, snake_alpha=None,
                 impl=None, block_n=0, split_out=None, a_ctx=0) -> None:
    """``split_out``: fp32 [S, M, N] -> split-K over S work units per tile; unit s writes its partial (gated) product to
    split_out[s]; the consumer (rmsnorm_mod / cfg_euler_tail ``delta=``) adds them to the residual stream."""
    d = _base_gemm(a, w, taps, batch, impl, block_n, a_ctx)
    d.epilogue, d.act = L.EPI_GENERIC, act
    M, N = a.shape[0], w.shape[0]
    if split_out is not None:
        _need(split_out, F32, "split_out")
        assert split_out.dim() == 3 and split_out.shape[1:] == (M, N) and split_out.is_contiguous()
        assert resid is None and out_f32 is None and out_bf16 is None and out2_bf16 is None and act == L.ACT_NONE
        d.k_splits, d.split_stride = split_out.shape[0], split_out.stride(0)
        out_f32 = split_out[0]
    if bias is not None:
        _need(bias, F32, "bias"); assert bias.numel() == N
        d.bias = bias.data_ptr()
    if gate is not None:
        _need(gate, F32, "gate"); assert gate.shape[-1] == N and gate.stride(-1) == 1
        d.gate = gate.data_ptr()
        d.gate_ld = gate.stride(0) if gate.dim() == 2 else 0
        d.rows_per_gate = rows_per_gate if rows_per_gate > 0 else (1 << 30)
    for t, name in ((resid, "resid"), (out_f32, "out_f32")):
        if t is not None:
            _need(t, F32, name); assert t.shape == (M, N) and t.stride(1) == 1
    if resid is not None:
        d.resid, d.ld_resid = resid.data_ptr(), resid.stride(0)
    if out_f32 is not None:
        d.out_f32, d.ld_out_f32 = out_f32.data_ptr(), out_f32.stride(0)
    if out_bf16 is not None:
        _need(out_bf16, BF16, "out_bf16"); assert out_bf16.shape == (M, N) and out_bf16.stride(1) == 1
        d.out_bf16, d.ld_out_bf16 = out_bf16.data_ptr(), out_bf16.stride(0)
    if out2_bf16 is not None:
        _need(out2_bf16, BF16, "out2_bf16"); _need(snake_alpha, F32, "snake_alpha")
        assert out2_bf16.shape == (M, N) and snake_alpha.numel() == N
        d.out2_bf16, d.ld_out2_bf16 = out2_bf16.data_ptr(), out2_bf16.stride(0)
        d.snake_alpha = snake_alpha.data_ptr()
    L.check(L.load().echo_gemm_bf16(C.byref(d), L.stream()))


def gemm_sections(a, w, sections: List[Tuple[int, torch.Tensor, Optional[torch.Tensor]]], *, head_dim: int,
                  rope_heads: int = 0, rope: Optional[Tuple[torch.Tensor, torch.Tensor]] = None,
                  pos_offset: int = 0, pos_stride: int = 1, rows_per_seq: int = 1, eps: float = 1e-5,
                  impl=None, block_n=0) -> None:
    """sections: [(SEC_type, out bf16 [M, width], norm_weight fp32 [width] or None)]; all same width."""
    d = _base_gemm(a, w, (0,), 1, impl, block_n)
    d.epilogue = L.EPI_SECTIONS
    M, N = a.shape[0], w.shape[0]
    ns = len(sections)
    width = N // ns
    d.n_sections, d.section_width, d.head_dim, d.rope_heads = ns, width, head_dim, rope_heads
    d.pos_offset, d.pos_stride, d.rows_per_seq, d.eps = pos_offset, pos_stride, rows_per_seq, eps
    if rope is not None:
        _need(rope[0], F32, "rope_cos"); _need(rope[1], F32, "rope_sin")
        assert rope[0].shape[1] == head_dim // 2 and rope[0].is_contiguous() and rope[1].is_contiguous()
        max_pos = pos_offset + (min(rows_per_seq, M) - 1) * pos_stride
        assert max_pos < rope[0].shape[0], "rope table too short"
        d.rope_cos, d.rope_sin = rope[0].data_ptr(), rope[1].data_ptr()
    ld = None
    for i, (typ, out, nw) in enumerate(sections):
        _need(out, BF16, "section.out"); assert out.shape == (M, width) and out.stride(1) == 1
        ld = out.stride(0) if ld is None else ld
        assert out.stride(0) == ld
        d.sections[i].type, d.sections[i].out = typ, out.data_ptr()
        if nw is not None:
            _need(nw, F32, "norm_weight"); assert nw.numel() == width and nw.is_contiguous()
            d.sections[i].norm_weight = nw.data_ptr()
    d.ld_sec_out = ld
    L.check(L.load().echo_gemm_bf16(C.byref(d), L.stream()))


def gemm_swiglu(a, w13, out_bf16, *, impl=None, block_n=0) -> None:
    """w13: rows interleaved in groups of 64 (w1 | w3), see pack_w13(); out [M, n/2] = silu(a w1^T) * (a w3^T)."""
    d = _base_gemm(a, w13, (0,), 1, impl, block_n)
    d.epilogue = L.EPI_SWIGLU
    _need(out_bf16, BF16, "out"); assert out_bf16.shape == (a.shape[0], w13.shape[0] // 2)
    d.out_bf16, d.ld_out_bf16 = out_bf16.data_ptr(), out_bf16.stride(0)
    L.check(L.load().echo_gemm_bf16(C.byref(d), L.stream()))


def pack_w13(w1: torch.Tensor, w3: torch.Tensor) -> torch.Tensor:
    """[F, K] x2 -> [2F, K] with 64-row groups alternating w1, w3 (F % 64 == 0)."""
    F_, K = w1.shape
    assert F_ % 64 == 0 and w3.shape == w1.shape
    return torch.stack([w1.reshape(F_ // 64, 64, K), w3.reshape(F_ // 64, 64, K)], dim=1).reshape(2 * F_, K).contiguous()


def sgemm(a, w, *, bias=None, addend=None, out=None, out_bf16=None, pre_silu=False, post=L.POST_NONE,
          out_scale: float = 1.0) -> None:
    """fp32 CUDA-core GEMM.  a [Z?, M, K], w [Z?, N, K], bias [Z?, N], addend [M, N] or [N] (shared over z)."""
    _need(a, F32, "sgemm.a"); _need(w, F32, "sgemm.w")
    d = L.SgemmDesc()
    nz = w.shape[0] if w.dim() == 3 else 1
    a3 = a if a.dim() == 3 else a.unsqueeze(0)
    w3 = w if w.dim() == 3 else w.unsqueeze(0)
    assert a3.stride(2) == 1 and w3.stride(2) == 1
    d.a, d.lda, d.a_zstride = a3.data_ptr(), a3.stride(1), (a3.stride(0) if a3.shape[0] > 1 else 0)
    d.w, d.ldw, d.w_zstride = w3.data_ptr(), w3.stride(1), (w3.stride(0) if nz > 1 else 0)
    d.m, d.k, d.n, d.nz = a3.shape[1], a3.shape[2], w3.shape[1], nz
    assert w3.shape[2] == d.k
    if bias is not None:
        _need(bias, F32, "sgemm.bias")
        b2 = bias if bias.dim() == 2 else bias.unsqueeze(0)
        assert b2.shape[1] == d.n and b2.stride(1) == 1
        d.bias, d.bias_zstride = b2.data_ptr(), (b2.stride(0) if b2.shape[0] > 1 else 0)
    if addend is not None:
        _need(addend, F32, "sgemm.addend"); assert addend.stride(-1) == 1 and addend.shape[-1] == d.n
        d.addend = addend.data_ptr()
        d.ld_addend = addend.stride(0) if addend.dim() == 2 else 0
    if out is not None:
        _need(out, F32, "sgemm.out")
        o3 = out if out.dim() == 3 else out.unsqueeze(0)
        assert o3.shape[1] == d.m and o3.shape[2] == d.n and o3.stride(2) == 1
        d.out, d.ldo, d.out_zstride = o3.data_ptr(), o3.stride(1), (o3.stride(0) if nz > 1 else 0)
    if out_bf16 is not None:
        _need(out_bf16, BF16, "sgemm.out_bf16")
        o3 = out_bf16 if out_bf16.dim() == 3 else out_bf16.unsqueeze(0)
        assert o3.shape[1] == d.m and o3.shape[2] == d.n and o3.stride(2) == 1
        d.out_bf16, d.ldo_bf16, d.out_bf16_zstride = o3.data_ptr(), o3.stride(1), (o3.stride(0) if nz > 1 else 0)
    d.pre_silu, d.post, d.out_scale = int(pre_silu), post, out_scale
    L.check(L.load().echo_sgemm_f32(C.byref(d), L.stream()))


def timestep_embedding(t: torch.Tensor, size: int) -> torch.Tensor:
    _need(t, F32, "t")
    out = torch.empty((t.numel(), size), device=t.device, dtype=F32)
    L.check(L.load().echo_timestep_embedding(t.data_ptr(), t.numel(), size, out.data_ptr(), L.stream()))
    return out


def rmsnorm_mod(x: torch.Tensor, scale1: torch.Tensor, shift: Optional[torch.Tensor], y: torch.Tensor, eps: float,
                rows_per_mod: int = 0, mod_ld: int = 0, step: Optional[torch.Tensor] = None,
                mod_step_stride: int = 0, delta: Optional[torch.Tensor] = None) -> None:
    """x fp32 [M, d] -> y bf16 [M, d]; scale1/shift point at table row 0 (row g at g*mod_ld).
    ``delta`` fp32 [S, M, d]: x += delta.sum(0) first (deferred residual update, x written back)."""
    _need(x, F32, "x"); _need(y, BF16, "y"); _need(scale1, F32, "scale1")
    M, d = x.shape
    if delta is not None:
        _need(delta, F32, "delta")
        assert delta.dim() == 3 and delta.shape[1:] == (M, d) and delta.is_contiguous() and x.stride(0) == d and step is None
        L.check(L.load().echo_add_rmsnorm_mod(x.data_ptr(), x.stride(0), M, d, eps, delta.data_ptr(), delta.shape[0],
                                              delta.stride(0), scale1.data_ptr(), L.ptr(shift), mod_ld, rows_per_mod,
                                              y.data_ptr(), y.stride(0), L.stream()))
        return
    L.check(L.load().echo_rmsnorm_mod(x.data_ptr(), x.stride(0), M, d, eps, scale1.data_ptr(), L.ptr(shift), mod_ld,
                                      rows_per_mod, L.ptr(step), mod_step_stride, y.data_ptr(), y.stride(0),
                                      L.stream()))


class Segment:
    """One KV segment for attention(): k/v bf16 [Bkv, L, H, Dh] (views allowed: strides are passed)."""

    def __init__(self, k, v, *, kv_batch_mod=0, mask=None, extent=None, kv_scale=None, causal=False, window=0,
                 length=None, q_offset=0):
        self.k, self.v, self.kv_batch_mod, self.mask, self.extent = k, v, kv_batch_mod, mask, extent
        self.kv_scale, self.causal, self.window = kv_scale, causal, window
        self.length = k.shape[1] if length is None else length
        self.q_offset = q_offset


ATTN_IMPL = 0     # echo_attn_desc.impl: 0 = tcgen05 kernel where eligible (head_dim 128, non-causal), else the general
                  # (causal / window / head_dim 64) kernel; tests pass impl=1 to run the general kernel at the DiT shapes


def attention(q: torch.Tensor, segs: Sequence[Segment], out: torch.Tensor, gate: Optional[torch.Tensor] = None,
              softmax_scale: Optional[float] = None, impl: Optional[int] = None, kv_splits: int = 0) -> None:
    """q bf16 [R, S, H, Dh] (last two dims contiguous); out bf16 [R*S, H*Dh].  ``kv_splits``: 0 = automatic split-KV
    (few query tiles: several CTAs share one tile's keys), 1 = off, 2..4 = forced."""
    _need(q, BF16, "q"); _need(out, BF16, "out")
    R, S, H, Dh = q.shape
    assert q.stride(3) == 1 and q.stride(2) == Dh and out.shape == (R * S, H * Dh) and out.is_contiguous()
    d = L.AttnDesc()
    d.q, d.q_row_stride, d.q_batch_stride = q.data_ptr(), q.stride(1), q.stride(0)
    d.out = out.data_ptr()
    if gate is not None:
        _need(gate, BF16, "gate"); assert gate.shape == out.shape and gate.is_contiguous()
        d.gate = gate.data_ptr()
    d.rows, d.seq, d.heads, d.head_dim = R, S, H, Dh
    d.softmax_scale = softmax_scale if softmax_scale is not None else Dh ** -0.5
    d.n_seg = len(segs)
    d.impl = ATTN_IMPL if impl is None else impl
    d.kv_splits = kv_splits
    if kv_splits != 1 and Dh == 128:
        ws = attn_workspace(q.device)
        d.workspace, d.workspace_bytes = ws.data_ptr(), ws.numel()
    for i, s in enumerate(segs):
        sg = d.seg[i]
        sg.len = s.length
        if s.length > 0:
            _need(s.k, BF16, "seg.k"); _need(s.v, BF16, "seg.v")
            assert s.k.shape[2:] == (H, Dh) and s.k.stride(3) == 1 and s.k.stride(2) == Dh
            assert s.v.stride() == s.k.stride()
            sg.k, sg.v = s.k.data_ptr(), s.v.data_ptr()
            sg.row_stride, sg.batch_stride = s.k.stride(1), s.k.stride(0)
        sg.kv_batch_mod = s.kv_batch_mod
        if s.mask is not None:
            assert s.mask.is_cuda and s.mask.dtype in (torch.bool, torch.uint8) and s.mask.shape[0] == R
            assert s.mask.stride(1) == 1 and s.mask.shape[1] >= s.length
            sg.mask, sg.mask_stride = s.mask.data_ptr(), s.mask.stride(0)
        if s.extent is not None:
            assert s.extent.dtype == torch.int32 and s.extent.numel() == R
            sg.extent = s.extent.data_ptr()
        if s.kv_scale is not None:
            _need(s.kv_scale, F32, "kv_scale")
            sg.kv_scale = s.kv_scale.data_ptr()
        sg.causal, sg.window, sg.q_offset = int(s.causal), s.window, s.q_offset
    L.check(L.load().echo_attention(C.byref(d), L.stream()))


def mask_extent(mask: torch.Tensor) -> torch.Tensor:
    assert mask.is_cuda and mask.dtype in (torch.bool, torch.uint8) and mask.dim() == 2 and mask.stride(1) == 1
    ext = torch.empty(mask.shape[0], device=mask.device, dtype=torch.int32)
    L.check(L.load().echo_mask_extent(mask.data_ptr(), mask.shape[0], mask.shape[1], mask.stride(0), ext.data_ptr(),
                                      L.stream()))
    return ext


def tail_desc(norm_w, w_out, b_out, *, positions, branches, eps, h=None, x_t=None, sched=None, step=None,
              cfg_text=0.0, cfg_speaker=0.0, rescale_k=None, rescale_sigma=None, v_out=None, v_comb_out=None,
              update_x=True, delta=None, push_to=None, guidance=None) -> L.TailDesc:
    """Descriptor of the fused sampler tail.  ``h`` (final residual stream) may be left out when the descriptor rides
    on ``echo_dit_step``, which fills it in.  ``push_to``: ((v_peer0, flag_peer0), (v_peer1, flag_peer1)) P2P-mapped
    tensors on the two other ranks of a CFG-branch split (branches == 1): the kernel stores its velocities there too
    and bumps the counters."""
    d = L.TailDesc()
    if push_to is not None:
        for k, (vp, fp) in enumerate(push_to):
            d.v_peer[k], d.flag_peer[k] = vp.data_ptr(), fp.data_ptr()
    for t, name in ((norm_w, "norm_w"), (w_out, "w_out"), (b_out, "b_out")):
        _need(t, F32, name); assert t.is_contiguous()
    d.norm_w, d.w_out, d.b_out = norm_w.data_ptr(), w_out.data_ptr(), b_out.data_ptr()
    d.positions, d.d, d.latent, d.branches = positions, w_out.shape[1], w_out.shape[0], branches
    if h is not None:
        _need(h, F32, "h")
        assert h.shape == (positions * branches, d.d)
        d.h, d.ldh = h.data_ptr(), h.stride(0)
        if delta is not None:
            _need(delta, F32, "delta")
            assert delta.dim() == 3 and delta.shape[1:] == h.shape and delta.is_contiguous() and h.is_contiguous()
            d.delta, d.n_delta, d.delta_stride = delta.data_ptr(), delta.shape[0], delta.stride(0)
    for t, name, n in ((x_t, "x_t", positions), (v_out, "v_out", positions * branches), (v_comb_out, "v_comb_out", positions)):
        if t is not None:
            _need(t, F32, name); assert t.is_contiguous() and t.numel() == n * d.latent
    if sched is not None:
        _need(sched, F32, "sched")
    d.x_t, d.v_out, d.v_comb_out = L.ptr(x_t), L.ptr(v_out), L.ptr(v_comb_out)
    d.sched, d.step = L.ptr(sched), L.ptr(step)
    d.eps, d.cfg_text, d.cfg_speaker = eps, cfg_text, cfg_speaker
    d.use_rescale = int(rescale_k is not None and rescale_sigma is not None)
    d.rescale_k, d.rescale_sigma = (rescale_k or 1.0), (rescale_sigma or 0.0)
    d.update_x = int(update_x)
    if guidance is not None:      # device fp32 [4] (cfg_text, cfg_speaker, rescale_k, rescale_sigma): overrides the immediates
        _need(guidance, F32, "guidance"); assert guidance.numel() == 4 and guidance.is_contiguous()
        d.guidance = guidance.data_ptr()
    return d


def cfg_euler_tail(h, norm_w, w_out, b_out, **kw) -> None:
    """Stand-alone launch of the fused tail on a given residual stream ``h`` (see tail_desc for the arguments)."""
    d = tail_desc(norm_w, w_out, b_out, h=h, **kw)
    L.check(L.load().echo_cfg_euler_tail(C.byref(d), L.stream()))


def cfg_combine_euler(v3: torch.Tensor, x_t: torch.Tensor, sched: torch.Tensor, *, cfg_text: float, cfg_speaker: float,
                      rescale_k=None, rescale_sigma=None, v_comb_out=None, wait=None) -> None:
    """v3 fp32 [3, positions, latent] (cond | text-dropped | speaker-dropped) -> x_t fp32 [positions, latent] += v dt.
    ``wait``: (counter0, counter1, value): spin until both int32 device counters (bumped by the peers' tail kernels)
    reach ``value`` before reading v3."""
    _need(v3, F32, "v3"); _need(x_t, F32, "x_t"); _need(sched, F32, "sched")
    assert v3.dim() == 3 and v3.shape[0] == 3 and v3.is_contiguous() and x_t.is_contiguous()
    d = L.TailDesc()
    d.v_out, d.x_t, d.v_comb_out, d.sched = v3.data_ptr(), x_t.data_ptr(), L.ptr(v_comb_out), sched.data_ptr()
    d.positions, d.latent, d.branches = v3.shape[1], v3.shape[2], 3
    assert x_t.numel() == v3.shape[1] * v3.shape[2]
    d.cfg_text, d.cfg_speaker = cfg_text, cfg_speaker
    d.use_rescale = int(rescale_k is not None and rescale_sigma is not None)
    d.rescale_k, d.rescale_sigma = (rescale_k or 1.0), (rescale_sigma or 0.0)
    d.update_x = 1
    if wait is not None:
        d.wait_flag[0], d.wait_flag[1], d.wait_value = wait[0].data_ptr(), wait[1].data_ptr(), int(wait[2])
    L.check(L.load().echo_cfg_combine_euler(C.byref(d), L.stream()))


def step_advance(step: torch.Tensor) -> None:
    L.check(L.load().echo_step_advance(step.data_ptr(), L.stream()))


def embedding(ids: torch.Tensor, table: torch.Tensor) -> torch.Tensor:
    assert ids.is_cuda and ids.dtype == torch.int32
    _need(table, F32, "table")
    out = torch.empty((ids.numel(), table.shape[1]), device=ids.device, dtype=F32)
    L.check(L.load().echo_embedding(ids.data_ptr(), ids.numel(), table.data_ptr(), table.shape[1], out.data_ptr(),
                                    L.stream()))
    return out


def dwconv_layernorm(x, w_dw, b_dw, ln_w, ln_b, eps, y, left_ctx: int = 0) -> None:
    """x fp32 [B, T, C] -> y bf16 [B, T, C].  ``left_ctx``: rows of the previous call's input that precede x in memory
    (streaming decode, B == 1)."""
    _need(x, F32, "x"); _need(y, BF16, "y")
    B, T, Cc = x.shape
    assert x.is_contiguous() and y.is_contiguous()
    L.check(L.load().echo_dwconv_layernorm_ctx(x.data_ptr(), B, T, Cc, left_ctx, w_dw.data_ptr(), b_dw.data_ptr(),
                                               ln_w.data_ptr(), ln_b.data_ptr(), eps, y.data_ptr(), L.stream()))


def conv_out_tanh(x_bf16, w, bias: float, wav, left_ctx: int = 0) -> None:
    """x bf16 [B, T, C] (Snake already applied), w fp32 [7, C] -> wav fp32 [B, T].  ``left_ctx`` as in dwconv_layernorm."""
    _need(x_bf16, BF16, "x"); _need(wav, F32, "wav"); _need(w, F32, "w")
    B, T, Cc = x_bf16.shape
    assert x_bf16.is_contiguous() and wav.is_contiguous() and w.shape == (7, Cc) and w.is_contiguous()
    L.check(L.load().echo_conv_out_tanh_ctx(x_bf16.data_ptr(), B, T, Cc, left_ctx, w.data_ptr(), bias, wav.data_ptr(),
                                            L.stream()))


def cast_snake(x, alpha, y) -> None:
    _need(x, F32, "x"); _need(y, BF16, "y")
    assert x.is_contiguous() and y.is_contiguous()
    Cc = x.shape[-1]
    L.check(L.load().echo_cast_snake(x.data_ptr(), x.numel() // Cc, Cc, L.ptr(alpha), y.data_ptr(), L.stream()))


def flattening_point(latent: torch.Tensor, window: int = 20, std_threshold: float = 0.05, target: float = 0.0) -> torch.Tensor:
    _need(latent, F32, "latent"); assert latent.dim() == 2 and latent.is_contiguous()
    out = torch.empty(1, device=latent.device, dtype=torch.int32)
    L.check(L.load().echo_flattening_point(latent.data_ptr(), latent.shape[0], latent.shape[1], window, std_threshold,
                                           target, out.data_ptr(), L.stream()))
    return out


def conv_in_snake(audio, w, bias, alpha, x, s) -> None:
    """audio fp32 [B, T] -> x fp32 [B, T, C] (causal conv k=7, w [C, 7]) and s bf16 [B, T, C] = snake(x, alpha)."""
    _need(audio, F32, "audio"); _need(x, F32, "x"); _need(s, BF16, "s"); _need(w, F32, "w")
    B, T = audio.shape
    Cc = w.shape[0]
    assert audio.is_contiguous() and x.is_contiguous() and s.is_contiguous() and w.shape == (Cc, 7) and w.is_contiguous()
    assert x.shape == (B, T, Cc) and s.shape == (B, T, Cc) and bias.numel() == Cc and alpha.numel() == Cc
    L.check(L.load().echo_conv_in_snake(audio.data_ptr(), B, T, Cc, w.data_ptr(), bias.data_ptr(), alpha.data_ptr(),
                                        x.data_ptr(), s.data_ptr(), L.stream()))


def snake_bf16(x, alpha, y) -> None:
    _need(x, BF16, "x"); _need(y, BF16, "y"); _need(alpha, F32, "alpha")
    Cc = x.shape[-1]
    assert x.is_contiguous() and y.is_contiguous() and y.shape == x.shape and alpha.numel() == Cc
    L.check(L.load().echo_snake_bf16(x.data_ptr(), x.numel() // Cc, Cc, alpha.data_ptr(), y.data_ptr(), L.stream()))


def rmsnorm_f32(x, weight, y, eps: float) -> None:
    """x fp32 [M, d] -> y fp32 [M, d] = x * rsqrt(mean x^2 + eps) * weight."""
    _need(x, F32, "x"); _need(y, F32, "y"); _need(weight, F32, "weight")
    M, d = x.shape
    assert y.shape == x.shape and x.stride(1) == 1 and y.stride(1) == 1 and weight.numel() == d
    L.check(L.load().echo_rmsnorm_f32(x.data_ptr(), x.stride(0), M, d, eps, weight.data_ptr(), y.data_ptr(), y.stride(0),
                                      L.stream()))


SPLIT_PLAIN, SPLIT_SNAKE, SPLIT_SWIGLU = 0, 1, 2


def split_bf16x3(x: torch.Tensor, out: Optional[torch.Tensor] = None, *, mode: int = SPLIT_PLAIN,
                 alpha: Optional[torch.Tensor] = None) -> torch.Tensor:
    """fp32 [M, C] (2C columns in SWIGLU mode) -> bf16 [M, 3C] = [hi | lo | hi]: the activation operand of a
    three-product split GEMM (csrc/precise.cu), with Snake / SwiGLU applied in fp32 first."""
    _need(x, F32, "x")
    assert x.dim() == 2 and x.stride(1) == 1
    M, Cc = x.shape
    if mode == SPLIT_SWIGLU:
        assert Cc % 2 == 0
        Cc //= 2
    if out is None:
        out = torch.empty((M, 3 * Cc), device=x.device, dtype=BF16)
    _need(out, BF16, "out")
    assert out.shape == (M, 3 * Cc) and out.stride(1) == 1
    if alpha is not None:
        _need(alpha, F32, "alpha"); assert alpha.numel() == Cc and alpha.is_contiguous()
    L.check(L.load().echo_split_bf16x3(x.data_ptr(), x.stride(0), M, Cc, mode, L.ptr(alpha), out.data_ptr(), out.stride(0),
                                       L.stream()))
    return out


def pack_split_weight(w2d: torch.Tensor, group: int) -> torch.Tensor:
    """fp32 [N, G * group] (G taps / sub-steps of `group` input channels each) -> bf16 [N, G * 3 * group] with every
    group laid out [w_hi | w_hi | w_lo]: the weight operand matching split_bf16x3's [hi | lo | hi] rows.  Load time."""
    N, K = w2d.shape
    assert K % group == 0
    w = w2d.float().reshape(N, K // group, group)
    hi = w.to(BF16)
    lo = (w - hi.float()).to(BF16)
    return torch.stack([hi, hi, lo], dim=2).reshape(N, 3 * K).contiguous()


def rope_f32(x: torch.Tensor, heads: int, head_dim: int, rope: Tuple[torch.Tensor, torch.Tensor], rows_per_seq: int,
             pos_offset: int = 0) -> None:
    """Rotary embedding in place on fp32 rows x[:, :heads * head_dim] (interleaved pairs, all heads)."""
    _need(x, F32, "x"); _need(rope[0], F32, "rope_cos"); _need(rope[1], F32, "rope_sin")
    assert x.dim() == 2 and x.stride(1) == 1 and x.shape[1] >= heads * head_dim
    assert rope[0].is_contiguous() and rope[1].is_contiguous() and rope[0].shape[1] == head_dim // 2
    L.check(L.load().echo_rope_f32(x.data_ptr(), x.stride(0), x.shape[0], heads, head_dim, rope[0].data_ptr(),
                                   rope[1].data_ptr(), rope[0].shape[0], rows_per_seq, pos_offset, L.stream()))


def attention_window_f32(q, k, v, out, *, batch: int, heads: int, head_dim: int, window: int) -> None:
    """fp32 causal sliding-window attention: q / k / v fp32 [B*T, >= H*Dh] views sharing one row stride, out fp32."""
    for t, n in ((q, "q"), (k, "k"), (v, "v"), (out, "out")):
        _need(t, F32, n); assert t.dim() == 2 and t.stride(1) == 1
    assert q.stride(0) == k.stride(0) == v.stride(0) and q.shape[0] % batch == 0 and out.shape[0] == q.shape[0]
    L.check(L.load().echo_attention_window_f32(q.data_ptr(), k.data_ptr(), v.data_ptr(), q.stride(0), batch,
                                               q.shape[0] // batch, heads, head_dim, window, head_dim ** -0.5,
                                               out.data_ptr(), out.stride(0), L.stream()))


def dwconv_layernorm_f32(x, w_dw, b_dw, ln_w, ln_b, eps, y) -> None:
    """x fp32 [B, T, C] -> y fp32 [B, T, C] (ConvNeXt depthwise conv k = 7 + LayerNorm, fp32 result)."""
    _need(x, F32, "x"); _need(y, F32, "y")
    B, T, Cc = x.shape
    assert x.is_contiguous() and y.is_contiguous()
    L.check(L.load().echo_dwconv_layernorm_f32(x.data_ptr(), B, T, Cc, w_dw.data_ptr(), b_dw.data_ptr(), ln_w.data_ptr(),
                                               ln_b.data_ptr(), eps, y.data_ptr(), L.stream()))


def vq_nearest(e, codebook_normalized, codebook, index_out, code_out=None) -> None:
    """e fp32 [M, 8] -> index_out int32 [M] (nearest l2-normalised code), code_out fp32 [M, 8] raw code vectors."""
    _need(e, F32, "e"); _need(codebook_normalized, F32, "codebook_normalized"); _need(codebook, F32, "codebook")
    M, dim = e.shape
    assert e.is_contiguous() and codebook.is_contiguous() and codebook_normalized.is_contiguous()
    assert codebook.shape == codebook_normalized.shape and codebook.shape[1] == dim
    assert index_out.is_cuda and index_out.dtype == torch.int32 and index_out.numel() == M and index_out.is_contiguous()
    if code_out is not None:
        _need(code_out, F32, "code_out"); assert code_out.shape == e.shape and code_out.is_contiguous()
    L.check(L.load().echo_vq_nearest(e.data_ptr(), M, dim, codebook_normalized.data_ptr(), codebook.data_ptr(),
                                     codebook.shape[0], index_out.data_ptr(), L.ptr(code_out), L.stream()))
