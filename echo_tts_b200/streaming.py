"""Streaming serving loop (SURVEY 8 f2 / BASELINE config 5): blockwise generation whose blocks are decoded to audio
as they are produced, without re-running anything that is already known.

The reference generates blockwise (``inference_blockwise.py:14-123``) but decodes only at the end and re-encodes the
whole latent prefix, three times over, at every block (:72-74); its README leaves streaming "not included".  Both
halves of the S1-DAC decoder and the latent-prefix encoder are causal, so the exact incremental forms exist:

* ``DecodeStream`` -- ``DAC.decode_zq`` (``autoencoder.py:1128-1132``) fed a few latent frames at a time.  State per
  layer: the last ``window - 1`` keys/values of each of the 8 post_module layers (``autoencoder.py:744-802``), and for
  every causal convolution the last ``(k - 1) * dilation`` input rows (``CausalConvNet`` left padding,
  ``autoencoder.py:283-289``; transposed convs keep one row).  A step lays each conv input out as
  ``[context rows | new rows]`` and points the implicit-GEMM at the new rows with ``a_ctx`` context rows in front
  (``echo_gemm_desc.a_ctx``), so the kernels are the ones the one-shot decode uses; an all-zero context is the
  reference's zero padding, hence stream == one-shot up to tile-order rounding.
* ``PrefixKVStream`` -- ``EchoDiT.get_kv_cache_latent`` (``model.py:623-636``) grown block by block: the 14 encoder
  layers are causal (``model.py:429-469``), so patches already encoded never change; a step encodes only the new
  patches against the cached per-layer encoder K/V (``echo_attn_segment.q_offset``) and appends their 24 (K, V)
  projections, RoPE'd at position 4 * patch, to one preallocated buffer.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, List, Optional, Sequence, Tuple

import torch

from . import _lib as L
from . import ops
from .autoencoder import DAC
from .model import EchoDiT, KVList

BF16, F32 = torch.bfloat16, torch.float32


class _Carry:
    """Named left-context rows carried from one step to the next.  ``load`` writes the carried rows (zeros on the first
    step) into buf[lead - n : lead]; ``save`` keeps the last n rows of buf[: lead + t] for the next step (when t < n
    those include the tail of the context just loaded)."""

    def __init__(self, device):
        self.device = device
        self.rows: Dict[str, torch.Tensor] = {}

    def load(self, name: str, buf: torch.Tensor, lead: int, n: int) -> None:
        st = self.rows.get(name)
        if st is None:
            buf[lead - n:lead].zero_()
        else:
            buf[lead - n:lead].copy_(st)

    def save(self, name: str, buf: torch.Tensor, lead: int, t: int, n: int) -> None:
        st = self.rows.get(name)
        if st is None:
            st = self.rows[name] = torch.empty((n,) + tuple(buf.shape[1:]), device=self.device, dtype=buf.dtype)
        st.copy_(buf[lead + t - n:lead + t])


class DecodeStream:
    """Incremental ``DAC.decode_zq`` for ONE utterance: ``step(z_q[1, D, t])`` returns the next ``t * hop`` samples.
    The concatenation of the steps' outputs equals ``decode_zq`` of the concatenated input (causality identity)."""

    LEAD = 54          # rows in front of the shared Snake buffer: the widest context, (7 - 1) * dilation 9

    def __init__(self, dac: DAC):
        self.dac = dac
        self.carry = _Carry(dac.device)
        self.pos = 0                                   # latent frames decoded so far (RoPE position, window fill)
        self.state = None                              # native path: every carried row in one device buffer
        if dac.use_native():
            lib, desc = L.load(), dac.native_desc()
            n = int(lib.echo_ae_decode_stream_state_bytes(C.byref(desc)))
            if n < 0:
                raise RuntimeError(f"echo_b200: {lib.echo_last_error().decode()}")
            self.state = torch.empty(n, device=dac.device, dtype=torch.uint8)
            L.check(lib.echo_ae_decode_stream_init(C.byref(desc), self.state.data_ptr(), n, L.stream()))

    def reset(self) -> None:
        """Start a new utterance on this stream object (state zeroed, position 0).  A serving loop keeps its streams --
        and the step graphs captured for them -- across requests (``DAC.stream_pool``)."""
        self.pos = 0
        self.carry.rows.clear()
        if self.state is not None:
            L.check(L.load().echo_ae_decode_stream_init(C.byref(self.dac.native_desc()), self.state.data_ptr(),
                                                        self.state.numel(), L.stream()))

    # -- post_module: 8 causal window-128 layers with a key/value cache ---------------------------------------------
    def _transformer(self, x: torch.Tensor, T: int) -> torch.Tensor:
        dac, cfg, dev = self.dac, self.dac.cfg, self.dac.device
        tw = dac.post_tr
        D = x.shape[1]
        H, Dh, W = cfg.tr_heads, cfg.tr_head_dim, cfg.tr_window
        if self.pos + T > tw["rope"][0].shape[0]:
            raise RuntimeError("stream longer than the post_module RoPE table (block_size, autoencoder.py:805-813)")
        C1 = W - 1                                     # cached keys a new query can still see
        nv = min(self.pos, C1)
        xn = torch.empty((T, D), device=dev, dtype=BF16)
        q = torch.empty((T, H * Dh), device=dev, dtype=BF16)
        kv = torch.empty((2, C1 + T, H * Dh), device=dev, dtype=BF16)
        ao = torch.empty((T, H * Dh), device=dev, dtype=BF16)
        hmid = torch.empty((T, cfg.tr_ffn), device=dev, dtype=BF16)
        for i, lw in enumerate(tw["layers"]):
            ops.rmsnorm_mod(x, lw["an"], None, xn, cfg.tr_eps)
            self.carry.load(f"tr{i}.k", kv[0], C1, C1)
            self.carry.load(f"tr{i}.v", kv[1], C1, C1)
            ops.gemm_sections(xn, lw["wqkv"], [(L.SEC_ROPE, q, None), (L.SEC_ROPE, kv[0, C1:], None),
                                               (L.SEC_PLAIN, kv[1, C1:], None)],
                              head_dim=Dh, rope_heads=H, rope=tw["rope"], pos_offset=self.pos, rows_per_seq=T,
                              eps=cfg.tr_eps)
            seg = ops.Segment(kv[0, C1 - nv:].view(1, nv + T, H, Dh), kv[1, C1 - nv:].view(1, nv + T, H, Dh),
                              causal=True, window=W, q_offset=nv)
            ops.attention(q.view(1, T, H, Dh), [seg], ao)
            self.carry.save(f"tr{i}.k", kv[0], C1, T, C1)
            self.carry.save(f"tr{i}.v", kv[1], C1, T, C1)
            ops.gemm_generic(ao, lw["wo"], gate=lw["ga"], resid=x, out_f32=x)
            ops.rmsnorm_mod(x, lw["fn"], None, xn, cfg.tr_eps)
            ops.gemm_swiglu(xn, lw["w13"], hmid)
            ops.gemm_generic(hmid, lw["w2"], gate=lw["gf"], resid=x, out_f32=x)
        ops.rmsnorm_mod(x, tw["norm"], None, xn, cfg.tr_eps)
        return xn

    # -- quantizer.upsample: transposed conv (k == stride: stateless) + ConvNeXt (depthwise k = 7: six rows) ----------
    def _upsample(self, a: torch.Tensor, T: int) -> Tuple[torch.Tensor, int]:
        """-> (buffer bf16 [6 + T', D] whose rows 6.. are the result, T')."""
        dac, dev = self.dac, self.dac.device
        D = dac.cfg.latent_dim
        out = None
        for u, up in enumerate(dac.ups):
            fac = up["factor"]
            xb = torch.empty((6 + T * fac, D), device=dev, dtype=F32)
            x = xb[6:]
            ops.gemm_generic(a, up["wt"], bias=up["bt"], out_f32=x.view(T, fac * D))
            T *= fac
            self.carry.load(f"up{u}.dw", xb, 6, 6)
            y = torch.empty((T, D), device=dev, dtype=BF16)
            ops.dwconv_layernorm(x.view(1, T, D), up["dw"], up["dwb"], up["lnw"], up["lnb"], 1e-6, y.view(1, T, D),
                                 left_ctx=6)
            self.carry.save(f"up{u}.dw", xb, 6, T, 6)
            hmid = torch.empty((T, up["pw1"].shape[0]), device=dev, dtype=BF16)
            ops.gemm_generic(y, up["pw1"], bias=up["pb1"], act=L.ACT_GELU, out_bf16=hmid)
            out = torch.empty((6 + T, D), device=dev, dtype=BF16)      # six rows in front: conv0's context
            a = out[6:]
            ops.gemm_generic(hmid, up["pw2"], bias=up["pb2"], gate=up["gamma"], resid=x, out_bf16=a)
        return out, T

    # -- Decoder: conv k=7, then per block transposed conv (one row) + three dilated residual units -------------------
    def _wave_decoder(self, ab: torch.Tensor, T: int) -> torch.Tensor:
        dac, dev, LEAD = self.dac, self.dac.device, self.LEAD
        C = dac.conv0_w.shape[0]
        self.carry.load("conv0", ab, 6, 6)
        sb = torch.empty((1 + T, C), device=dev, dtype=BF16)           # one row in front: the transposed conv's context
        ops.gemm_generic(ab[6:], dac.conv0_w, taps=tuple(-(6 - k) for k in range(7)), a_ctx=6, bias=dac.conv0_b,
                         out2_bf16=sb[1:], snake_alpha=dac.blocks[0]["alpha_in"])
        self.carry.save("conv0", ab, 6, T, 6)
        lead = 1
        for bi, blk in enumerate(dac.blocks):
            r = blk["rate"]
            Cout = blk["wt"].shape[0] // r
            units = blk["units"]
            x = torch.empty((T, r * Cout), device=dev, dtype=F32)
            s1b = torch.empty((LEAD + T * r, Cout), device=dev, dtype=BF16)
            s1 = s1b[LEAD:]
            self.carry.load(f"b{bi}.up", sb, lead, 1)
            ops.gemm_generic(sb[lead:], blk["wt"], taps=(0, -1), a_ctx=1, bias=blk["bt"], out_f32=x,
                             out2_bf16=s1.view(T, r * Cout), snake_alpha=units[0]["a1"].repeat(r).contiguous())
            self.carry.save(f"b{bi}.up", sb, lead, T, 1)
            T *= r
            x = x.view(T, Cout)
            s2 = torch.empty((T, Cout), device=dev, dtype=BF16)
            for j, u in enumerate(units):
                d = u["dil"]
                n = 6 * d
                self.carry.load(f"b{bi}.u{j}", s1b, LEAD, n)
                ops.gemm_generic(s1, u["w7"], taps=tuple(-(6 - k) * d for k in range(7)), a_ctx=n, bias=u["b7"],
                                 out2_bf16=s2, snake_alpha=u["a2"])
                self.carry.save(f"b{bi}.u{j}", s1b, LEAD, T, n)
                if j + 1 < len(units):
                    nxt = units[j + 1]["a1"]
                elif bi + 1 < len(dac.blocks):
                    nxt = dac.blocks[bi + 1]["alpha_in"]
                else:
                    nxt = dac.alpha_out
                ops.gemm_generic(s2, u["w1"], bias=u["b1"], resid=x, out2_bf16=s1, snake_alpha=nxt,
                             out_f32=(x if j + 1 < len(units) else None))   # after the last unit only snake(x) is read
            sb, lead = s1b, LEAD
        wav = torch.empty((1, T), device=dev, dtype=F32)
        self.carry.load("out", sb, lead, 6)
        ops.conv_out_tanh(sb[lead:].view(1, T, -1), dac.w_last, dac.b_last, wav, left_ctx=6)
        self.carry.save("out", sb, lead, T, 6)
        return wav

    def _run(self, x: torch.Tensor, T: int) -> torch.Tensor:
        if self.state is not None:       # one native call (csrc/ae_decode.cu: echo_ae_decode_stream_step)
            dac = self.dac
            if self.pos + T > dac.post_tr["rope"][0].shape[0]:
                raise RuntimeError("stream longer than the post_module RoPE table (block_size, autoencoder.py:805-813)")
            ws = dac.native_workspace(1, T, True)
            wav = torch.empty((1, T * dac.cfg.hop), device=dac.device, dtype=F32)
            L.check(L.load().echo_ae_decode_stream_step(C.byref(dac.native_desc()), self.state.data_ptr(), x.data_ptr(), T,
                                                        self.pos, wav.data_ptr(), ws.data_ptr(), ws.numel(), L.stream()))
            self.pos += T
            return wav.unsqueeze(1)
        a = self._transformer(x, T)
        ab, T2 = self._upsample(a, T)
        wav = self._wave_decoder(ab, T2)
        self.pos += T
        return wav.unsqueeze(1)

    @torch.inference_mode()
    def step(self, z_q: torch.Tensor) -> torch.Tensor:
        """z_q [1, latent_dim, t] (the next t frames) -> waveform [1, 1, t * hop]."""
        B, D, T = z_q.shape
        if B != 1:
            raise ValueError("a DecodeStream carries the state of one utterance: open one stream per batch row")
        x = z_q.to(self.dac.device, F32).transpose(1, 2).contiguous()
        if x.data_ptr() == z_q.data_ptr():       # the decode clobbers x
            x = x.clone()
        x = x.view(T, D)
        return self._run(x, T).to(self.dac.dtype)

    @torch.inference_mode()
    def step_latent(self, z: torch.Tensor, pca_components: torch.Tensor, pca_mean: torch.Tensor,
                    latent_scale: float) -> torch.Tensor:
        """``ae_decode`` (inference.py:232-235) of the next frames: z fp32 [1, t, 80] -> waveform fp32 [1, 1, t * hop]."""
        B, T, lat = z.shape
        if B != 1:
            raise ValueError("a DecodeStream carries the state of one utterance: open one stream per batch row")
        dev = self.dac.device
        w, mean = self.dac.pca_operands(pca_components, pca_mean)

        def launch(zin: torch.Tensor) -> torch.Tensor:
            x = torch.empty((T, w.shape[0]), device=dev, dtype=F32)
            ops.sgemm(zin.view(T, lat), w, addend=mean, out=x, out_scale=1.0 / latent_scale)
            return self._run(x, T)

        zin = z.to(dev, F32).contiguous()
        from .sampler import _use_graphs
        if self.state is None or not _use_graphs():
            return launch(zin)
        # CUDA graph per (frames, stream position): a step is ~230 launches of a few microseconds each (launch-bound when
        # issued one by one).  The first step at a key runs plainly (it also sizes the cached scratch), the second is
        # captured, later ones replay; the carried state lives in this stream's one state buffer, so replays continue it.
        graphs = self.__dict__.setdefault("_step_graphs", {})
        key = (T, self.pos, float(latent_scale), w.data_ptr())
        ent = graphs.setdefault(key, dict(calls=0, graph=None))
        ent["calls"] += 1
        if ent["calls"] == 1:
            return launch(zin)
        if ent["graph"] is None:
            ent["z"] = zin.clone()
            pos0 = self.pos
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                ent["wav"] = launch(ent["z"])
            self.pos = pos0                    # the capture advanced the position without running anything
            ent["graph"] = g
        ent["z"].copy_(zin)
        ent["graph"].replay()
        self.pos += T
        return ent["wav"].clone()


class PrefixKVStream:
    """``EchoDiT.get_kv_cache_latent`` (model.py:623-636) grown incrementally for a batch of B prompts.

    ``append(latent[B, t, 80])`` encodes the patches completed by the new frames (frames that do not yet fill a patch
    wait for the next call) and appends their per-decoder-layer (K, V) to ``self.kv`` -- a ``KVList`` over one buffer
    laid out [layers, 2, B, capacity, 2048], of which the first ``self.patches`` keys per row are valid."""

    def __init__(self, model: EchoDiT, batch: int, capacity_frames: int):
        if not model.has_blockwise:
            raise RuntimeError("blockwise modules were deleted at load (delete_blockwise_modules=True)")
        cfg, dev = model.cfg, model.device
        self.model, self.B = model, batch
        p = cfg.speaker_patch_size
        self.cap = max(1, capacity_frames // p)
        D = cfg.model_size
        self.buf = torch.zeros((cfg.num_layers, 2, batch, self.cap, D), device=dev, dtype=BF16)
        self.kv = KVList((self.buf[l, 0].view(batch, self.cap, cfg.num_heads, 128),
                          self.buf[l, 1].view(batch, self.cap, cfg.num_heads, 128)) for l in range(cfg.num_layers))
        self.kv.buf = self.buf
        d_enc = model.lat_in_w.shape[0]
        n_enc = len(model.lat_enc.layers)
        self.enc_kv = torch.zeros((n_enc, 2, batch, self.cap, d_enc), device=dev, dtype=BF16)   # encoder-layer K / V
        self.patches = 0
        self.pending = torch.empty((batch, 0, cfg.latent_size), device=dev, dtype=F32)

    @torch.inference_mode()
    def append(self, latent: torch.Tensor) -> int:
        """-> number of valid prefix patches after the call."""
        model, cfg, dev, B = self.model, self.model.cfg, self.model.device, self.B
        p = cfg.speaker_patch_size
        lat = torch.cat([self.pending, latent.to(dev, F32)], 1)
        n_new = lat.shape[1] // p
        self.pending = lat[:, n_new * p:].contiguous()
        if n_new == 0:
            return self.patches
        P0 = self.patches
        if P0 + n_new > self.cap:
            raise RuntimeError("PrefixKVStream: capacity exceeded")
        enc = model.lat_enc
        H = enc.heads
        d = model.lat_in_w.shape[0]
        Dm, Hm = cfg.model_size, cfg.num_heads
        if model.use_native_encoders():          # one native call; it walks the prompt rows itself (csrc/encode_kv.cu)
            xin = lat[:, :n_new * p].contiguous().view(B * n_new, p * cfg.latent_size)
            x = torch.empty((B * n_new, d), device=dev, dtype=F32)
            ops.sgemm(xin, model.lat_in_w, bias=model.lat_in_b, out=x, out_scale=1.0 / 6.0)          # model.py:461-462
            model.encode_kv_native(enc, x, B, n_new, causal=True, mask=None, norm_w=model.lat_norm, key="kv_latent",
                                   rope_patch=True, past=P0, enc_cache=self.enc_kv, kv_out=self.buf)
            self.patches = P0 + n_new
            return self.patches
        # one pass per prompt row: rows of a batch are independent, and a single-row view keeps [cached | new] keys
        # contiguous in the preallocated buffers
        for b in range(B):
            xin = lat[b, :n_new * p].contiguous().view(n_new, p * cfg.latent_size)
            x = torch.empty((n_new, d), device=dev, dtype=F32)
            ops.sgemm(xin, model.lat_in_w, bias=model.lat_in_b, out=x, out_scale=1.0 / 6.0)        # model.py:461-462
            xn = torch.empty((n_new, d), device=dev, dtype=BF16)
            q = torch.empty((n_new, d), device=dev, dtype=BF16)
            g = torch.empty((n_new, d), device=dev, dtype=BF16)
            ao = torch.empty((n_new, d), device=dev, dtype=BF16)
            hmid = torch.empty((n_new, enc.layers[0]["w2"].shape[1]), device=dev, dtype=BF16)
            for i, lw in enumerate(enc.layers):
                kc, vc = self.enc_kv[i, 0, b], self.enc_kv[i, 1, b]                               # [cap, d]
                ops.rmsnorm_mod(x, lw["an"], None, xn, cfg.norm_eps)
                ops.gemm_sections(xn, lw["qkvg"],
                                  [(L.SEC_NORM_ROPE, q, lw["qn"]), (L.SEC_NORM_ROPE, kc[P0:P0 + n_new], lw["kn"]),
                                   (L.SEC_PLAIN, vc[P0:P0 + n_new], None), (L.SEC_SIGMOID, g, None)],
                                  head_dim=128, rope_heads=H, rope=model.rope, pos_offset=P0, rows_per_seq=n_new,
                                  eps=cfg.norm_eps)
                seg = ops.Segment(kc[:P0 + n_new].view(1, P0 + n_new, H, 128), vc[:P0 + n_new].view(1, P0 + n_new, H, 128),
                                  causal=True, q_offset=P0)
                ops.attention(q.view(1, n_new, H, 128), [seg], ao, gate=g)
                ops.gemm_generic(ao, lw["wo"], resid=x, out_f32=x)
                ops.rmsnorm_mod(x, lw["mn"], None, xn, cfg.norm_eps)
                ops.gemm_swiglu(xn, lw["w13"], hmid)
                ops.gemm_generic(hmid, lw["w2"], resid=x, out_f32=x)
            sn = torch.empty((n_new, d), device=dev, dtype=BF16)
            ops.rmsnorm_mod(x, model.lat_norm, None, sn, cfg.norm_eps)
            for l, lw in enumerate(model.layers):                                                   # model.py:284-293
                ops.gemm_sections(sn, lw["kv_latent"],
                                  [(L.SEC_NORM_ROPE, self.buf[l, 0, b, P0:P0 + n_new], lw["kn"]),
                                   (L.SEC_PLAIN, self.buf[l, 1, b, P0:P0 + n_new], None)],
                                  head_dim=128, rope_heads=Hm // 2, rope=model.rope, pos_offset=P0 * p, pos_stride=p,
                                  rows_per_seq=n_new, eps=cfg.norm_eps)
        self.patches = P0 + n_new
        return self.patches
