"""EchoDiT engine: the reference's model object (``model.py:472-642``) re-built on the sm_100a kernels.

Same duck-typed surface the reference samplers use (``.device``, ``.dtype``, ``__call__``,
``get_kv_cache_text/speaker/latent``), but nothing here is an ``nn.Module``: weights are repacked once
at load (wq|wk|wv|gate fused to one N=4d operand, w1/w3 interleaved for the SwiGLU epilogue, per-layer
K|V projections of the conditioning fused, everything GEMM-shaped in bf16, everything t-only or tiny
kept fp32), and a forward is ~8 fused kernel launches per layer.

Precision policy: fp32 residual stream, fp32 timestep / AdaLN tables / in_proj / out_proj,
bf16 tensor-core operands with fp32 accumulation, fp32 softmax statistics.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, List, Optional, Tuple

import torch

from . import _lib as L
from . import ops
from .schema import DiTConfig, ECHO_BASE

KVCache = List[Tuple[torch.Tensor, torch.Tensor]]
BF16, F32 = torch.bfloat16, torch.float32


class KVList(list):
    """List[(K, V)] per decoder layer, all views of one buffer ``buf`` bf16 [layers, 2, B*L, D]."""
    buf: torch.Tensor

    @staticmethod
    def from_buffer(buf: torch.Tensor, B: int, Ls: int, H: int) -> "KVList":
        out = KVList((buf[l, 0].view(B, Ls, H, 128), buf[l, 1].view(B, Ls, H, 128)) for l in range(buf.shape[0]))
        out.buf = buf
        return out


def rope_table(head_dim: int, n_pos: int, device) -> Tuple[torch.Tensor, torch.Tensor]:
    """(cos, sin) fp32 [n_pos, head_dim/2], computed on the host exactly as the reference does
    (model.py:9-14) and uploaded once instead of on every forward (model.py:579)."""
    inv = 1.0 / (10000.0 ** (torch.arange(0, head_dim, 2)[: head_dim // 2] / head_dim))
    ang = torch.outer(torch.arange(n_pos), inv)
    return torch.cos(ang).to(device).contiguous(), torch.sin(ang).to(device).contiguous()


class _EncoderWeights:
    """Packed weights of a 14-layer conditioning encoder (TextEncoder / SpeakerEncoder, model.py:392-469)."""

    def __init__(self, sd, prefix: str, n_layers: int, heads: int, dev):
        self.heads = heads
        self.layers = []
        for i in range(n_layers):
            p = f"{prefix}.blocks.{i}"
            a = f"{p}.attention"
            g = lambda k: sd[k].to(dev)
            self.layers.append(dict(
                qkvg=torch.cat([g(f"{a}.wq.weight"), g(f"{a}.wk.weight"), g(f"{a}.wv.weight"),
                                g(f"{a}.gate.weight")], 0).to(BF16).contiguous(),
                qn=g(f"{a}.q_norm.weight").float().reshape(-1).contiguous(),
                kn=g(f"{a}.k_norm.weight").float().reshape(-1).contiguous(),
                wo=g(f"{a}.wo.weight").to(BF16).contiguous(),
                w13=ops.pack_w13(g(f"{p}.mlp.w1.weight"), g(f"{p}.mlp.w3.weight")).to(BF16),
                w2=g(f"{p}.mlp.w2.weight").to(BF16).contiguous(),
                an=g(f"{p}.attention_norm.weight").float().contiguous(),
                mn=g(f"{p}.mlp_norm.weight").float().contiguous(),
            ))


class EchoDiT:
    """Drop-in for the reference ``EchoDiT`` on the sampling path (inference only)."""

    def __init__(self, state_dict: Dict[str, torch.Tensor], cfg: DiTConfig = ECHO_BASE, device="cuda",
                 dtype: torch.dtype = torch.bfloat16, max_positions: int = 8192):
        L.load()
        if not torch.cuda.is_available():
            raise RuntimeError("echo_tts_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
        if dtype != torch.bfloat16:
            raise ValueError("echo_tts_b200 computes GEMMs in bf16 (fp32 accumulate); pass dtype=torch.bfloat16")
        assert cfg.head_dim == 128, "kernels are specialised for head_dim 128"
        self.cfg = cfg
        self._device = torch.device(device)
        self._dtype = dtype
        dev = self._device
        sd = state_dict
        g = lambda k: sd[k].to(dev)
        D, Lr = cfg.model_size, cfg.num_layers
        self.has_blockwise = "latent_norm.weight" in sd

        # fp32 small / t-only weights
        self.cond_w = [g(f"cond_module.{i}.weight").float().contiguous() for i in (0, 2, 4)]
        self.in_w, self.in_b = g("in_proj.weight").float().contiguous(), g("in_proj.bias").float().contiguous()
        self.out_norm = g("out_norm.weight").float().contiguous()
        self.out_w, self.out_b = g("out_proj.weight").float().contiguous(), g("out_proj.bias").float().contiguous()
        parts = ("shift", "scale", "gate")
        adas = ("attention_adaln", "mlp_adaln")
        self.ada_down = [torch.cat([g(f"blocks.{l}.{a}.{p}_down.weight").float() for l in range(Lr) for a in adas], 0)
                         .contiguous() for p in parts]                      # 3 x [L*2*r, D]
        self.ada_up = [torch.stack([g(f"blocks.{l}.{a}.{p}_up.weight").float() for l in range(Lr) for a in adas], 0)
                       .contiguous() for p in parts]                        # 3 x [L*2, D, r]
        self.ada_up_b = [torch.stack([g(f"blocks.{l}.{a}.{p}_up.bias").float() for l in range(Lr) for a in adas], 0)
                         .contiguous() for p in parts]                      # 3 x [L*2, D]

        # decoder layers (bf16 tensor-core operands)
        self.layers = []
        for l in range(Lr):
            a = f"blocks.{l}.attention"
            lw = dict(
                qkvg=torch.cat([g(f"{a}.wq.weight"), g(f"{a}.wk.weight"), g(f"{a}.wv.weight"), g(f"{a}.gate.weight")],
                               0).to(BF16).contiguous(),
                qn=g(f"{a}.q_norm.weight").float().reshape(-1).contiguous(),
                kn=g(f"{a}.k_norm.weight").float().reshape(-1).contiguous(),
                wo=g(f"{a}.wo.weight").to(BF16).contiguous(),
                w13=ops.pack_w13(g(f"blocks.{l}.mlp.w1.weight"), g(f"blocks.{l}.mlp.w3.weight")).to(BF16),
                w2=g(f"blocks.{l}.mlp.w2.weight").to(BF16).contiguous(),
                kv_text=torch.cat([g(f"{a}.wk_text.weight"), g(f"{a}.wv_text.weight")], 0).to(BF16).contiguous(),
                kv_speaker=torch.cat([g(f"{a}.wk_speaker.weight"), g(f"{a}.wv_speaker.weight")], 0).to(BF16).contiguous(),
            )
            if self.has_blockwise:
                lw["kv_latent"] = torch.cat([g(f"{a}.wk_latent.weight"), g(f"{a}.wv_latent.weight")], 0).to(BF16).contiguous()
            self.layers.append(lw)

        # conditioning encoders
        self.text_emb = g("text_encoder.text_embedding.weight").float().contiguous()
        self.text_enc = _EncoderWeights(sd, "text_encoder", cfg.text_num_layers, cfg.text_num_heads, dev)
        self.text_norm = g("text_norm.weight").float().contiguous()
        self.spk_enc = _EncoderWeights(sd, "speaker_encoder", cfg.speaker_num_layers, cfg.speaker_num_heads, dev)
        self.spk_in_w = g("speaker_encoder.in_proj.weight").float().contiguous()
        self.spk_in_b = g("speaker_encoder.in_proj.bias").float().contiguous()
        self.spk_norm = g("speaker_norm.weight").float().contiguous()
        if self.has_blockwise:
            self.lat_enc = _EncoderWeights(sd, "latent_encoder", cfg.speaker_num_layers, cfg.speaker_num_heads, dev)
            self.lat_in_w = g("latent_encoder.in_proj.weight").float().contiguous()
            self.lat_in_b = g("latent_encoder.in_proj.bias").float().contiguous()
            self.lat_norm = g("latent_norm.weight").float().contiguous()

        self.rope = rope_table(cfg.head_dim, max_positions, dev)
        self._ws: Dict[Tuple[int, int], dict] = {}
        self._ones = torch.ones(1, device=dev, dtype=F32)

    # ------------------------------------------------------------------------------------------
    @property
    def device(self) -> torch.device:
        return self._device

    @property
    def dtype(self) -> torch.dtype:
        return self._dtype

    def eval(self):
        return self

    # ------------------------------------------------------------------------------------------
    # t-only work: timestep embedding -> cond_module -> all 48 LowRankAdaLN modulations, once per
    # schedule instead of once per forward (model.py:583, 70-74, 81).
    # ------------------------------------------------------------------------------------------
    def build_mod_table(self, t: torch.Tensor) -> torch.Tensor:
        """t fp32 [T] -> table fp32 [T, L, 2, 3, D]: (shift, 1 + scale, tanh(gate)) for (attention, mlp)."""
        cfg = self.cfg
        D, Lr, r = cfg.model_size, cfg.num_layers, cfg.adaln_rank
        T = t.numel()
        dev = self._device
        t = t.to(dev, F32).contiguous()
        if getattr(self, "round_t_to_model_dtype", False) and self._dtype != F32:
            # the reference's default bf16 path rounds t to the model dtype before the sinusoidal embedding
            # (inference.py:517 `.to(dtype)`, model.py:39); off by default: the parity oracle is the fp32 path
            t = t.to(self._dtype).to(F32)
        emb = ops.timestep_embedding(t, cfg.timestep_embed_size)
        h1 = torch.empty((T, D), device=dev, dtype=F32)
        h2 = torch.empty((T, D), device=dev, dtype=F32)
        cond = torch.empty((T, 3 * D), device=dev, dtype=F32)
        ops.sgemm(emb, self.cond_w[0], out=h1, post=L.POST_SILU)
        ops.sgemm(h1, self.cond_w[1], out=h2, post=L.POST_SILU)
        ops.sgemm(h2, self.cond_w[2], out=cond)
        table = torch.empty((T, Lr, 2, 3, D), device=dev, dtype=F32)
        tv = table.view(T, Lr * 2, 3, D)
        lo = torch.empty((T, Lr * 2 * r), device=dev, dtype=F32)
        for p, post in enumerate((L.POST_NONE, L.POST_PLUS1, L.POST_TANH)):
            c_p = cond[:, p * D:(p + 1) * D]
            ops.sgemm(c_p, self.ada_down[p], out=lo, pre_silu=True)
            ops.sgemm(lo.view(T, Lr * 2, r).permute(1, 0, 2), self.ada_up[p], bias=self.ada_up_b[p], addend=c_p,
                      out=tv[:, :, p, :].permute(1, 0, 2), post=post)
        return table

    # ------------------------------------------------------------------------------------------
    # conditioning encoders (once per prompt)
    # ------------------------------------------------------------------------------------------
    def _run_encoder(self, enc: _EncoderWeights, x: torch.Tensor, B: int, Ls: int, *, causal: bool,
                     mask: Optional[torch.Tensor]) -> torch.Tensor:
        """x fp32 [B*Ls, d] residual stream, updated in place (EncoderTransformerBlock, model.py:335-339)."""
        cfg = self.cfg
        d = x.shape[1]
        H = enc.heads
        M = B * Ls
        dev = self._device
        xn = torch.empty((M, d), device=dev, dtype=BF16)
        qkvg = torch.empty((4, M, d), device=dev, dtype=BF16)
        ao = torch.empty((M, d), device=dev, dtype=BF16)
        hmid = torch.empty((M, enc.layers[0]["w2"].shape[1]), device=dev, dtype=BF16)
        ext = ops.mask_extent(mask) if mask is not None else None
        for lw in enc.layers:
            ops.rmsnorm_mod(x, lw["an"], None, xn, cfg.norm_eps)
            ops.gemm_sections(xn, lw["qkvg"],
                              [(L.SEC_NORM_ROPE, qkvg[0], lw["qn"]), (L.SEC_NORM_ROPE, qkvg[1], lw["kn"]),
                               (L.SEC_PLAIN, qkvg[2], None), (L.SEC_SIGMOID, qkvg[3], None)],
                              head_dim=128, rope_heads=H, rope=self.rope, rows_per_seq=Ls, eps=cfg.norm_eps)
            seg = ops.Segment(qkvg[1].view(B, Ls, H, 128), qkvg[2].view(B, Ls, H, 128), mask=mask, extent=ext,
                              causal=causal)
            ops.attention(qkvg[0].view(B, Ls, H, 128), [seg], ao, gate=qkvg[3])
            ops.gemm_generic(ao, lw["wo"], resid=x, out_f32=x)
            ops.rmsnorm_mod(x, lw["mn"], None, xn, cfg.norm_eps)
            ops.gemm_swiglu(xn, lw["w13"], hmid)
            ops.gemm_generic(hmid, lw["w2"], resid=x, out_f32=x)
        return x

    def _project_kv(self, state: torch.Tensor, norm_w: torch.Tensor, key: str, B: int, Ls: int,
                    rope_patch: bool = False) -> KVCache:
        """state fp32 [B*Ls, d_enc] -> per decoder layer (K, V) bf16 [B, Ls, H, 128]  (model.py:270-293)."""
        cfg = self.cfg
        D, H = cfg.model_size, cfg.num_heads
        M = B * Ls
        sn = torch.empty((M, state.shape[1]), device=self._device, dtype=BF16)
        ops.rmsnorm_mod(state, norm_w, None, sn, cfg.norm_eps)
        buf = torch.empty((cfg.num_layers, 2, M, D), device=self._device, dtype=BF16)
        for l, lw in enumerate(self.layers):
            ops.gemm_sections(sn, lw[key], [(L.SEC_NORM_ROPE, buf[l, 0], lw["kn"]), (L.SEC_PLAIN, buf[l, 1], None)],
                              head_dim=128, rope_heads=(H // 2 if rope_patch else 0), rope=self.rope,
                              pos_stride=(cfg.speaker_patch_size if rope_patch else 1), rows_per_seq=Ls,
                              eps=cfg.norm_eps)
        return KVList.from_buffer(buf, B, Ls, H)

    # ------------------------------------------------------------------------------------------
    # native form of _run_encoder + _project_kv: ONE call (csrc/encode_kv.cu: echo_encode_kv); ECHO_ENC_NATIVE=0 keeps
    # the Python launch sequence above (A/B, bit-identical)
    # ------------------------------------------------------------------------------------------
    @staticmethod
    def use_native_encoders() -> bool:
        return os.environ.get("ECHO_ENC_NATIVE", "1") != "0"

    def _enc_arrays(self, enc: _EncoderWeights, key: str):
        """HOST arrays of weight pointers for echo_encode_kv (built once per encoder / projection family)."""
        cache = self.__dict__.setdefault("_c_enc", {})
        ent = cache.get((id(enc), key))
        if ent is None:
            la = (L.EncLayer * len(enc.layers))()
            for i, lw in enumerate(enc.layers):
                la[i].w_qkvg, la[i].w_o, la[i].w13, la[i].w2 = (lw[k].data_ptr() for k in ("qkvg", "wo", "w13", "w2"))
                la[i].q_norm, la[i].k_norm = lw["qn"].data_ptr(), lw["kn"].data_ptr()
                la[i].attn_norm, la[i].mlp_norm = lw["an"].data_ptr(), lw["mn"].data_ptr()
            wkv = (C.c_void_p * len(self.layers))(*[lw[key].data_ptr() for lw in self.layers])
            kn = (C.c_void_p * len(self.layers))(*[lw["kn"].data_ptr() for lw in self.layers])
            ent = cache[(id(enc), key)] = (la, wkv, kn)
        return ent

    def encode_kv_native(self, enc: _EncoderWeights, x: torch.Tensor, B: int, Ls: int, *, causal: bool,
                         mask: Optional[torch.Tensor], norm_w: torch.Tensor, key: str, rope_patch: bool = False,
                         past: int = 0, enc_cache: Optional[torch.Tensor] = None, kv_out: Optional[torch.Tensor] = None
                         ) -> torch.Tensor:
        """x fp32 [B*Ls, d_enc] (clobbered) -> kv_out bf16 [layers, 2, B, cap, D] with positions [past, past+Ls) written.
        ``enc_cache`` bf16 [enc_layers, 2, B, cap_e, d_enc] holds the encoder's own K/V across incremental calls."""
        cfg, dev = self.cfg, self._device
        la, wkv, kn = self._enc_arrays(enc, key)
        if kv_out is None:
            kv_out = torch.empty((cfg.num_layers, 2, B, Ls, cfg.model_size), device=dev, dtype=BF16)
        ext = ops.mask_extent(mask) if mask is not None else None
        d = L.EncodeKVDesc()
        d.n_layers, d.d, d.heads, d.ffn = len(enc.layers), x.shape[1], enc.heads, enc.layers[0]["w2"].shape[1]
        d.eps, d.causal, d.layers, d.out_norm = cfg.norm_eps, int(causal), la, norm_w.data_ptr()
        d.rope_cos, d.rope_sin, d.rope_len = self.rope[0].data_ptr(), self.rope[1].data_ptr(), self.rope[0].shape[0]
        d.rows, d.seq, d.past, d.x = B, Ls, past, x.data_ptr()
        if mask is not None:
            assert mask.dtype in (torch.bool, torch.uint8) and mask.stride(1) == 1 and mask.shape[0] == B
            d.mask, d.mask_stride, d.extent = mask.data_ptr(), mask.stride(0), ext.data_ptr()
        if enc_cache is not None:
            assert enc_cache.is_contiguous() and enc_cache.dtype == BF16 and enc_cache.shape[:3] == (len(enc.layers), 2, B)
            d.enc_cache, d.enc_cap = enc_cache.data_ptr(), enc_cache.shape[3]
        assert kv_out.is_contiguous() and kv_out.dtype == BF16 and kv_out.shape[:3] == (cfg.num_layers, 2, B)
        d.dec_layers, d.dec_d, d.dec_heads = cfg.num_layers, cfg.model_size, cfg.num_heads
        d.w_kv, d.k_norm = C.cast(wkv, C.POINTER(C.c_void_p)), C.cast(kn, C.POINTER(C.c_void_p))
        d.rope_heads, d.pos_stride = (cfg.num_heads // 2, cfg.speaker_patch_size) if rope_patch else (0, 1)
        d.kv_out, d.out_cap = kv_out.data_ptr(), kv_out.shape[3]
        gws = ops.gemm_workspace(dev)
        d.gemm_impl, d.gemm_workspace, d.gemm_workspace_bytes = ops.GEMM_IMPL, gws.data_ptr(), gws.numel()
        lib = L.load()
        n = int(lib.echo_encode_kv_workspace_bytes(C.byref(d)))
        if n < 0:
            raise RuntimeError(f"echo_b200: {lib.echo_last_error().decode()}")
        ws = torch.empty(n, device=dev, dtype=torch.uint8)
        L.check(lib.echo_encode_kv(C.byref(d), ws.data_ptr(), n, L.stream()))
        return kv_out

    def _encode(self, enc, x, B, Ls, *, causal, mask, norm_w, key, rope_patch=False) -> KVCache:
        if self.use_native_encoders():
            buf = self.encode_kv_native(enc, x, B, Ls, causal=causal, mask=mask, norm_w=norm_w, key=key, rope_patch=rope_patch)
            return KVList.from_buffer(buf.view(self.cfg.num_layers, 2, B * Ls, self.cfg.model_size), B, Ls, self.cfg.num_heads)
        self._run_encoder(enc, x, B, Ls, causal=causal, mask=mask)
        return self._project_kv(x, norm_w, key, B, Ls, rope_patch=rope_patch)

    @torch.inference_mode()
    def kv_caches_text_speaker(self, text_input_ids, text_mask, speaker_latent) -> Tuple[KVCache, KVCache]:
        """``get_kv_cache_text`` and ``get_kv_cache_speaker`` (inference.py:482-483) side by side: the two encoders are
        independent stacks of small launches (14 layers at 768 tokens / a few dozen patches), so the speaker encoder runs on
        a second stream forked from and joined back into the current one (inside a CUDA-graph capture this becomes two
        parallel branches of the graph).  The side stream always waits for the current stream first, so buffers the caching
        allocator hands back to it are never reused under work the current stream still has in flight."""
        if ops.GEMM_IMPL == 3 or os.environ.get("ECHO_ENCODER_STREAMS") == "0":
            # (the stream-K schedule exchanges partial tiles through ONE workspace per device: no concurrent GEMM streams)
            return self.get_kv_cache_text(text_input_ids, text_mask), self.get_kv_cache_speaker(speaker_latent)
        cur = torch.cuda.current_stream(self._device)
        side = self.__dict__.get("_side_stream")
        if side is None:
            side = self._side_stream = torch.cuda.Stream(device=self._device)
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            ops.ATTN_WS_SLOT = 1                 # its attention launches may overlap the text encoder's: own exchange buffer
            try:
                kv_spk = self.get_kv_cache_speaker(speaker_latent)
            finally:
                ops.ATTN_WS_SLOT = 0
        kv_text = self.get_kv_cache_text(text_input_ids, text_mask)
        cur.wait_stream(side)
        return kv_text, kv_spk

    def get_kv_cache_text(self, text_input_ids: torch.Tensor, text_mask: Optional[torch.Tensor]) -> KVCache:
        """model.py:606-613."""
        B, Lt = text_input_ids.shape
        ids = text_input_ids.to(self._device, torch.int32).contiguous().view(-1)
        x = ops.embedding(ids, self.text_emb)
        mask = text_mask.to(self._device).contiguous() if text_mask is not None else None
        return self._encode(self.text_enc, x, B, Lt, causal=False, mask=mask, norm_w=self.text_norm, key="kv_text")

    def _patch_encode(self, enc, in_w, in_b, latent: torch.Tensor):
        cfg = self.cfg
        B, T, Cc = latent.shape
        p = cfg.speaker_patch_size
        assert T % p == 0, "latent length must be a multiple of the patch size (model.py:459)"
        Ls = T // p
        xin = latent.to(self._device, F32).contiguous().view(B * Ls, Cc * p)
        x = torch.empty((B * Ls, in_w.shape[0]), device=self._device, dtype=F32)
        ops.sgemm(xin, in_w, bias=in_b, out=x, out_scale=1.0 / 6.0)          # model.py:461-462
        return x, B, Ls

    @torch.inference_mode()
    def get_kv_cache_speaker(self, speaker_latent: torch.Tensor) -> KVCache:
        """model.py:615-621."""
        x, B, Ls = self._patch_encode(self.spk_enc, self.spk_in_w, self.spk_in_b, speaker_latent)
        return self._encode(self.spk_enc, x, B, Ls, causal=True, mask=None, norm_w=self.spk_norm, key="kv_speaker")

    @torch.inference_mode()
    def get_kv_cache_latent(self, prefix_latent: torch.Tensor) -> KVCache:
        """model.py:623-636 (+ half-head RoPE at positions 4*patch, model.py:284-293)."""
        if not self.has_blockwise:
            raise RuntimeError("blockwise modules were deleted at load (delete_blockwise_modules=True)")
        if prefix_latent.shape[1] == 0:
            z = torch.empty((prefix_latent.shape[0], 0, self.cfg.num_heads, 128), device=self._device, dtype=BF16)
            return [(z, z) for _ in range(self.cfg.num_layers)]
        x, B, Ls = self._patch_encode(self.lat_enc, self.lat_in_w, self.lat_in_b, prefix_latent)
        return self._encode(self.lat_enc, x, B, Ls, causal=True, mask=None, norm_w=self.lat_norm, key="kv_latent",
                            rope_patch=True)

    # ------------------------------------------------------------------------------------------
    # one velocity evaluation (EchoDiT.forward, model.py:563-604) -- ends at the final residual stream;
    # out_norm/out_proj live in the fused sampler tail.
    # ------------------------------------------------------------------------------------------
    WS_CAPACITY = 8      # activation workspaces kept per model (LRU); captured graphs hold their own references

    def workspace(self, rows: int, S: int) -> dict:
        key = (rows, S)
        ws = self._ws.get(key)
        if ws is not None:
            self._ws[key] = self._ws.pop(key)        # most recently used last (dicts keep insertion order)
        if ws is None:
            while len(self._ws) >= self.WS_CAPACITY:
                self._ws.pop(next(iter(self._ws)))
            cfg, dev = self.cfg, self._device
            M, D = rows * S, cfg.model_size
            ns = self.k_splits(M)
            ws = dict(x=torch.empty((M, D), device=dev, dtype=F32), xn=torch.empty((M, D), device=dev, dtype=BF16),
                      delta=(torch.empty((ns, M, D), device=dev, dtype=F32) if ns > 1 else None),
                      qkvg=torch.empty((4, M, D), device=dev, dtype=BF16), ao=torch.empty((M, D), device=dev, dtype=BF16),
                      h=torch.empty((M, cfg.intermediate_size), device=dev, dtype=BF16))
            self._ws[key] = ws
        return ws

    def k_splits(self, M: int) -> int:
        """Split-K factor for the two N = model_size projections (wo, w2) of a forward over M tokens: with few tokens
        (one un-guided row of 640) a 256 x 256 tiling leaves most of the 74 CTA pairs idle, so the K loop is split
        over enough work units to fill them; the partial products are summed by the next norm kernel."""
        tiles = -(-M // 256) * -(-self.cfg.model_size // 256)
        cap = 8 if M <= 256 else 4        # one row tile (a 160-latent streaming block): w2's 92 k-blocks over 8 units
        return max(1, min(cap, 74 // tiles, self.cfg.model_size // 64))

    def _layer_array(self):
        """HOST array of per-layer weight pointers for echo_dit_step (built once; the tensors live in self.layers)."""
        arr = self.__dict__.get("_c_layers")
        if arr is None:
            arr = (L.DitLayer * len(self.layers))()
            for l, lw in enumerate(self.layers):
                arr[l].w_qkvg, arr[l].w_o = lw["qkvg"].data_ptr(), lw["wo"].data_ptr()
                arr[l].w13, arr[l].w2 = lw["w13"].data_ptr(), lw["w2"].data_ptr()
                arr[l].q_norm, arr[l].k_norm = lw["qn"].data_ptr(), lw["kn"].data_ptr()
            self._c_layers = arr
        return arr

    def _fill_kv(self, dst: "L.DitKV", cache: KVCache, R: int, B: int, keep: list, *, length=None, mask=None,
                 extent=None, kv_scale=None) -> None:
        cfg = self.cfg
        H, Dh = cfg.num_heads, cfg.head_dim
        assert len(cache) == cfg.num_layers
        k0 = cache[0][0]
        dst.len = k0.shape[1] if length is None else length
        if dst.len > 0:
            for k, v in cache:
                for t in (k, v):
                    if not t.is_cuda or t.dtype != BF16:
                        raise RuntimeError("kv cache: bf16 CUDA tensors required (no CPU fallback)")
                    assert t.shape == k0.shape and t.stride() == k0.stride()
            assert k0.shape[2:] == (H, Dh) and k0.stride(3) == 1 and k0.stride(2) == Dh and dst.len <= k0.shape[1]
            ka = (C.c_void_p * len(cache))(*[k.data_ptr() for k, _ in cache])
            va = (C.c_void_p * len(cache))(*[v.data_ptr() for _, v in cache])
            keep += [ka, va]
            dst.k, dst.v = C.cast(ka, C.POINTER(C.c_void_p)), C.cast(va, C.POINTER(C.c_void_p))
            dst.row_stride, dst.batch_stride = k0.stride(1), k0.stride(0)
        dst.kv_batch_mod = B if k0.shape[0] != R else 0
        if mask is not None:
            assert mask.is_cuda and mask.dtype in (torch.bool, torch.uint8) and mask.shape[0] >= R
            assert mask.stride(1) == 1 and mask.shape[1] >= dst.len
            dst.mask, dst.mask_stride = mask.data_ptr(), mask.stride(0)
        if extent is not None:
            assert extent.is_cuda and extent.dtype == torch.int32 and extent.numel() >= R and extent.is_contiguous()
            dst.extent = extent.data_ptr()
        if kv_scale is not None:
            assert kv_scale.is_cuda and kv_scale.dtype == F32 and kv_scale.numel() >= cfg.num_layers and kv_scale.is_contiguous()
            dst.kv_scale = kv_scale.data_ptr()

    def step_desc(self, x_in: torch.Tensor, mod: torch.Tensor, B: int, R: int, *, text_mask, text_extent, spk_mask,
                  spk_extent, kv_text: KVCache, kv_speaker: KVCache, kv_latent: Optional[KVCache] = None,
                  start_pos: int = 0, n_prefix: int = 0, spk_scale: Optional[torch.Tensor] = None,
                  tail: Optional["L.TailDesc"] = None):
        """Descriptor of one velocity evaluation for ``echo_dit_step`` (+ the objects it points into, to be kept alive
        by the caller until the launch calls have returned).

        x_in: fp32 [Bx, S, latent] with Bx == R (one input per row) or R % Bx == 0 (CFG: the three branches share x_t,
        inference.py:516 -- in_proj is evaluated once per prompt and written to every branch).
        mod: fp32 [R or 1, L, 2, 3, D] modulation rows (row g serves rows g*S..); text_mask/spk_mask uint8/bool
        [>= R, L_t] / [>= R, L_s] (patch rate) with their extents; spk_scale: fp32 [L] per-layer speaker K/V multiplier
        (speaker_kv_scale, inference.py:420-428) or None; ``tail``: ops.tail_desc(...) to fuse out_norm / out_proj /
        CFG combine / Euler behind the forward."""
        cfg = self.cfg
        for t, name in ((x_in, "x_in"), (mod, "mod")):
            if not t.is_cuda or t.dtype != F32:
                raise RuntimeError(f"{name}: fp32 CUDA tensor required (no CPU fallback)")
        Bx, S, lat = x_in.shape
        assert x_in.is_contiguous() and lat == cfg.latent_size and R % Bx == 0
        assert mod.dim() == 5 and mod.shape[1:] == (cfg.num_layers, 2, 3, cfg.model_size) and mod[0].is_contiguous()
        assert mod.shape[0] in (1, R)
        ws = self.workspace(R, S)
        keep = [x_in, mod, ws, tail, text_mask, text_extent, spk_mask, spk_extent, spk_scale, kv_text, kv_speaker, kv_latent]
        d = L.DitStepDesc()
        d.n_layers, d.d, d.heads, d.head_dim = cfg.num_layers, cfg.model_size, cfg.num_heads, cfg.head_dim
        d.ffn, d.latent, d.eps = cfg.intermediate_size, cfg.latent_size, cfg.norm_eps
        d.layers = self._layer_array()
        d.in_w, d.in_b = self.in_w.data_ptr(), self.in_b.data_ptr()
        d.rope_cos, d.rope_sin, d.rope_len = self.rope[0].data_ptr(), self.rope[1].data_ptr(), self.rope[0].shape[0]
        d.x_in, d.x_rows, d.rows, d.seq, d.pos_offset = x_in.data_ptr(), Bx, R, S, start_pos
        d.mod, d.mod_row_stride, d.mod_rows = mod.data_ptr(), mod.stride(0), mod.shape[0]
        d.k_splits = self.k_splits(R * S)
        if kv_latent is not None and n_prefix > 0:
            self._fill_kv(d.prefix, kv_latent, R, B, keep, length=n_prefix)
        self._fill_kv(d.text, kv_text, R, B, keep, mask=text_mask, extent=text_extent)
        self._fill_kv(d.speaker, kv_speaker, R, B, keep, mask=spk_mask, extent=spk_extent, kv_scale=spk_scale)
        d.x, d.xn, d.qkvg, d.ao, d.hmid = (ws[k].data_ptr() for k in ("x", "xn", "qkvg", "ao", "h"))
        d.delta = L.ptr(ws["delta"])
        gws = ops.gemm_workspace(self._device)
        d.gemm_workspace, d.gemm_workspace_bytes = gws.data_ptr(), gws.numel()
        aws = ops.attn_workspace(self._device)
        d.attn_workspace, d.attn_workspace_bytes = aws.data_ptr(), aws.numel()
        d.gemm_impl, d.attn_impl = ops.GEMM_IMPL, ops.ATTN_IMPL
        if tail is not None:
            d.tail = C.pointer(tail)
        return d, keep

    def forward_hidden(self, x_in: torch.Tensor, mod: torch.Tensor, B: int, *, text_mask, **kw):
        """x_in fp32 [Bx, S, latent] -> (final residual stream fp32 [R*S, D], pending delta or None): when the forward
        runs split-K the last w2's partial products [splits, R*S, D] are still to be added to the stream (the sampler
        tail does it).  One native call (``echo_dit_step``); arguments as in step_desc."""
        R = text_mask.shape[0]
        d, keep = self.step_desc(x_in, mod, B, R, text_mask=text_mask, **kw)
        L.check(L.load().echo_dit_step(C.byref(d), L.stream()))
        ws = self.workspace(R, x_in.shape[1])
        return ws["x"], ws["delta"]

    @torch.inference_mode()
    def __call__(self, x, t, text_mask, speaker_mask, kv_cache_text, kv_cache_speaker, start_pos=None,
                 kv_cache_latent=None) -> torch.Tensor:
        """Reference-compatible forward (model.py:563-604): velocities fp32 [R, S, latent].
        Third-party samplers written against the reference call this; our own sampler bypasses it."""
        cfg, dev = self.cfg, self._device
        start_pos = start_pos or 0
        R, S, lat = x.shape
        mod = self.build_mod_table(t.to(dev).float().reshape(-1))
        tm = text_mask.to(dev).contiguous()
        sm = speaker_mask.to(dev)[..., ::cfg.speaker_patch_size].contiguous()
        n_prefix = 0
        if kv_cache_latent is not None and kv_cache_latent[0][0].shape[1] > 0:
            n_prefix = min(kv_cache_latent[0][0].shape[1], -(-start_pos // cfg.speaker_patch_size))
        Bkv = kv_cache_text[0][0].shape[0]
        h, pending = self.forward_hidden(x.to(dev).float().contiguous(), mod, Bkv, text_mask=tm, text_extent=ops.mask_extent(tm),
                                spk_mask=sm, spk_extent=ops.mask_extent(sm), kv_text=kv_cache_text,
                                kv_speaker=kv_cache_speaker, kv_latent=kv_cache_latent, start_pos=start_pos,
                                n_prefix=n_prefix)
        v = torch.empty((R * S, lat), device=dev, dtype=F32)
        ops.cfg_euler_tail(h, self.out_norm, self.out_w, self.out_b, positions=R * S, branches=1, eps=cfg.norm_eps,
                           v_out=v, update_x=False, delta=pending)
        return v.view(R, S, lat)

    forward = __call__
