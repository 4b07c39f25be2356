"""Fish S1-DAC engine on sm_100a kernels: ``DAC.decode_zq`` (reference ``autoencoder.py:1128-1132``) and, when the
checkpoint carries the encoder / quantizer tensors, ``DAC.encode_zq`` (``autoencoder.py:1080-1126``).

Activations are time-major ``[B, T, C]`` so that every convolution is an implicit GEMM over TMA-shifted row
windows (causal left padding = TMA out-of-bounds zero fill), weight-norm is folded once at load
(``autoencoder.py:90-94`` recomputes it every forward), Snake / bias / residual / GELU / layer-scale live in GEMM
epilogues, and the dense [T, T] window mask built on the CPU per call (``autoencoder.py:762-797``) is replaced by
in-kernel window arithmetic.  bf16 tensor-core operands, fp32 accumulation, fp32 residual streams.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, Optional

import torch

from . import _lib as L
from . import ops
from .schema import AEConfig, AE_BASE

BF16, F32 = torch.bfloat16, torch.float32


def _fold_wn(sd, prefix: str, dev) -> torch.Tensor:
    g = sd[f"{prefix}.conv.parametrizations.weight.original0"].to(dev).float()
    v = sd[f"{prefix}.conv.parametrizations.weight.original1"].to(dev).float()
    return g * v / v.flatten(1).norm(dim=1).reshape(-1, *([1] * (v.dim() - 1)))


def _pack_conv(w: torch.Tensor, dtype=BF16) -> torch.Tensor:
    """Conv1d weight [Cout, Cin, K] -> GEMM B operand [Cout, K*Cin] (tap-major K axis), bf16."""
    return w.permute(0, 2, 1).reshape(w.shape[0], -1).to(dtype).contiguous()


def _fold_wn_plain(sd, prefix: str, dev) -> torch.Tensor:
    """weight-normed conv without the CausalConvNet wrapper (VectorQuantize.in_proj / out_proj, autoencoder.py:126-127)."""
    g = sd[f"{prefix}.parametrizations.weight.original0"].to(dev).float()
    v = sd[f"{prefix}.parametrizations.weight.original1"].to(dev).float()
    return g * v / v.flatten(1).norm(dim=1).reshape(-1, *([1] * (v.dim() - 1)))


def _pack_conv_strided(w: torch.Tensor, stride: int, dtype=BF16) -> torch.Tensor:
    """Strided causal Conv1d weight [Cout, Cin, K], K = taps*stride (left pad K - stride, autoencoder.py:283-289) ->
    GEMM B operand [Cout, taps * stride*Cin] over the input viewed as [T/stride, stride*Cin] rows:
    out[t] = sum_tap sum_j sum_ci x[(t - (taps-1) + tap)*stride + j, ci] * w[co, ci, tap*stride + j]."""
    cout, cin, k = w.shape
    taps = k // stride
    assert taps * stride == k
    return w.reshape(cout, cin, taps, stride).permute(0, 2, 3, 1).reshape(cout, k * cin).to(dtype).contiguous()


def _pack_convT(w: torch.Tensor, stride: int) -> torch.Tensor:
    """ConvTranspose1d weight [Cin, Cout, K] (K = stride or 2*stride) -> [stride*Cout, taps*Cin]:
    out[t*stride + j] = sum_tap x[t - tap] . w[:, :, j + tap*stride]   (autoencoder.py:310-316, right trim)."""
    cin, cout, k = w.shape
    taps = k // stride
    wt = w.reshape(cin, cout, taps, stride).permute(3, 1, 2, 0)       # [j, co, tap, ci]
    return wt.reshape(stride * cout, taps * cin).to(BF16).contiguous()


class DAC:
    """Decode half of the reference ``DAC`` (drop-in for ``fish_ae.decode_zq`` / ``.dtype`` / ``.device``)."""

    def __init__(self, state_dict: Dict[str, torch.Tensor], cfg: AEConfig = AE_BASE, device="cuda",
                 dtype: torch.dtype = torch.float32):
        L.load()
        if not torch.cuda.is_available():
            raise RuntimeError("echo_tts_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
        self.cfg = cfg
        self._device = torch.device(device)
        self._dtype = dtype                     # interface dtype of decode_zq's input/output (reference default fp32)
        dev, sd = self._device, state_dict
        f = lambda k: sd[k].to(dev).float().contiguous()
        self.post_tr = self._load_transformer(sd, "quantizer.post_module", cfg.tr_layers)
        self.ups = []
        for u, fac in enumerate(cfg.up_factors):
            p = f"quantizer.upsample.{u}"
            self.ups.append(dict(
                factor=fac, wt=_pack_convT(sd[f"{p}.0.conv.weight"].to(dev).float(), fac),
                bt=f(f"{p}.0.conv.bias").repeat(fac).contiguous(),
                dw=f(f"{p}.1.dwconv.conv.weight").reshape(-1, 7).contiguous(), dwb=f(f"{p}.1.dwconv.conv.bias"),
                lnw=f(f"{p}.1.norm.weight"), lnb=f(f"{p}.1.norm.bias"),
                pw1=sd[f"{p}.1.pwconv1.weight"].to(dev).to(BF16).contiguous(), pb1=f(f"{p}.1.pwconv1.bias"),
                pw2=sd[f"{p}.1.pwconv2.weight"].to(dev).to(BF16).contiguous(), pb2=f(f"{p}.1.pwconv2.bias"),
                gamma=f(f"{p}.1.gamma")))
        dm = "decoder.model"
        self.conv0_w, self.conv0_b = _pack_conv(_fold_wn(sd, f"{dm}.0", dev)), f(f"{dm}.0.conv.bias")
        self.blocks = []
        for i, r in enumerate(cfg.decoder_rates):
            b = f"{dm}.{i + 1}.block"
            blk = dict(rate=r, alpha_in=f(f"{b}.0.alpha").reshape(-1),
                       wt=_pack_convT(_fold_wn(sd, f"{b}.1", dev), r), bt=f(f"{b}.1.conv.bias").repeat(r).contiguous(),
                       units=[])
            for j, dil in enumerate((1, 3, 9)):
                ru = f"{b}.{2 + j}.block"
                blk["units"].append(dict(dil=dil, a1=f(f"{ru}.0.alpha").reshape(-1), a2=f(f"{ru}.2.alpha").reshape(-1),
                                         w7=_pack_conv(_fold_wn(sd, f"{ru}.1", dev)), b7=f(f"{ru}.1.conv.bias"),
                                         w1=_pack_conv(_fold_wn(sd, f"{ru}.3", dev)), b1=f(f"{ru}.3.conv.bias")))
            self.blocks.append(blk)
        n = len(cfg.decoder_rates)
        self.alpha_out = f(f"{dm}.{n + 1}.alpha").reshape(-1)
        w_last = _fold_wn(sd, f"{dm}.{n + 2}", dev)                       # [1, C, 7]
        self.w_last = w_last[0].t().contiguous()                         # [7, C] fp32
        self.b_last = float(sd[f"{dm}.{n + 2}.conv.bias"].float().item())
        self._pca_cache = None
        self.has_encoder = "encoder.block.0.conv.bias" in sd
        if self.has_encoder:
            self._load_encoder(sd)

    def _load_transformer(self, sd, pm: str, n_layers: int) -> dict:
        """Weights of one WindowLimitedTransformer (autoencoder.py:744-802); its RoPE table is the module's own
        bf16-rounded buffer (autoencoder.py:805-813)."""
        dev = self._device
        f = lambda k: sd[k].to(dev).float().contiguous()
        fc = sd[f"{pm}.freqs_cis"].to(dev).float()
        layers = []
        for i in range(n_layers):
            p = f"{pm}.layers.{i}"
            layers.append(dict(
                wqkv=sd[f"{p}.attention.wqkv.weight"].to(dev).to(BF16).contiguous(),
                wo=sd[f"{p}.attention.wo.weight"].to(dev).to(BF16).contiguous(),
                w13=ops.pack_w13(sd[f"{p}.feed_forward.w1.weight"].to(dev), sd[f"{p}.feed_forward.w3.weight"].to(dev)).to(BF16),
                w2=sd[f"{p}.feed_forward.w2.weight"].to(dev).to(BF16).contiguous(),
                an=f(f"{p}.attention_norm.weight"), fn=f(f"{p}.ffn_norm.weight"),
                ga=f(f"{p}.attention_layer_scale.gamma"), gf=f(f"{p}.ffn_layer_scale.gamma")))
        return dict(layers=layers, norm=f(f"{pm}.norm.weight"), rope=(fc[..., 0].contiguous(), fc[..., 1].contiguous()))

    def _load_convnext(self, sd, p: str) -> dict:
        dev = self._device
        f = lambda k: sd[k].to(dev).float().contiguous()
        return dict(dw=f(f"{p}.dwconv.conv.weight").reshape(-1, 7).contiguous(), dwb=f(f"{p}.dwconv.conv.bias"),
                    lnw=f(f"{p}.norm.weight"), lnb=f(f"{p}.norm.bias"),
                    pw1=sd[f"{p}.pwconv1.weight"].to(dev).to(BF16).contiguous(), pb1=f(f"{p}.pwconv1.bias"),
                    pw2=sd[f"{p}.pwconv2.weight"].to(dev).to(BF16).contiguous(), pb2=f(f"{p}.pwconv2.bias"),
                    gamma=f(f"{p}.gamma"))

    def _load_encoder(self, sd) -> None:
        """Encoder (autoencoder.py:903-929), quantizer.downsample / pre_module (:416-425, :1150-1163) and the
        1 + n_codebooks VectorQuantize layers (:117-157), packed like the decode half."""
        cfg, dev = self.cfg, self._device
        f = lambda k: sd[k].to(dev).float().contiguous()
        eb = "encoder.block"
        self.enc_in_w = _fold_wn(sd, f"{eb}.0", dev)[:, 0, :].contiguous()          # [C, 7] fp32
        self.enc_in_b = f(f"{eb}.0.conv.bias")
        self.enc_blocks = []
        n = len(cfg.encoder_rates)
        for i, r in enumerate(cfg.encoder_rates):
            b = f"{eb}.{i + 1}.block"
            blk = dict(rate=r, units=[], alpha_out=f(f"{b}.3.alpha").reshape(-1), bd=f(f"{b}.4.conv.bias"))
            for j, dil in enumerate((1, 3, 9)):
                ru = f"{b}.{j}.block"
                blk["units"].append(dict(dil=dil, a1=f(f"{ru}.0.alpha").reshape(-1), a2=f(f"{ru}.2.alpha").reshape(-1),
                                         b7=f(f"{ru}.1.conv.bias"), b1=f(f"{ru}.3.conv.bias")))
            self.enc_blocks.append(blk)
        self.enc_tr = (self._load_transformer(sd, f"{eb}.{n}.block.5", cfg.enc_tr_layers) if cfg.enc_tr_layers > 0 else None)
        self.enc_alpha_out = f(f"{eb}.{n + 1}.alpha").reshape(-1)
        self.enc_out_b = f(f"{eb}.{n + 2}.conv.bias")
        self.downs = []
        for u, fac in enumerate(cfg.up_factors):
            p = f"quantizer.downsample.{u}"
            d = self._load_convnext(sd, f"{p}.1")
            d.update(factor=fac, bd=f(f"{p}.0.conv.bias"))
            del d["pw1"], d["pw2"]
            self.downs.append(d)
        self.pre_tr = self._load_transformer(sd, "quantizer.pre_module", cfg.tr_layers)
        self.quantizers = []
        names = ["quantizer.semantic_quantizer.quantizers.0"] + [f"quantizer.quantizer.quantizers.{q}" for q in range(cfg.n_codebooks)]
        for p in names:
            w_in = _fold_wn_plain(sd, f"{p}.in_proj", dev)[:, :, 0].contiguous()          # [dim, D] fp32
            w_out = _fold_wn_plain(sd, f"{p}.out_proj", dev)[:, :, 0].contiguous()        # [D, dim] fp32
            cb = f(f"{p}.codebook.weight")
            self.quantizers.append(dict(w_in=w_in, b_in=f(f"{p}.in_proj.bias"), w_out=w_out, b_out=f(f"{p}.out_proj.bias"),
                                        w_out_neg=(-w_out).contiguous(), b_out_neg=(-f(f"{p}.out_proj.bias")).contiguous(),
                                        cb=cb, cbn=torch.nn.functional.normalize(cb).contiguous()))   # autoencoder.py:149
        self._load_encoder_split(sd)

    def _load_encoder_split(self, sd) -> None:
        """The encode half's GEMM weights once more as two-term bf16 splits ([w_hi | w_hi | w_lo] per group of input
        channels, ``ops.pack_split_weight``): operands of the three-product fp32-class GEMMs in front of the RVQ's
        integer decision (csrc/precise.cu)."""
        cfg, dev = self.cfg, self._device
        eb = "encoder.block"
        n = len(cfg.encoder_rates)
        C = cfg.encoder_dim
        for i, blk in enumerate(self.enc_blocks):
            b = f"{eb}.{i + 1}.block"
            for j, u in enumerate(blk["units"]):
                ru = f"{b}.{j}.block"
                u["w7s"] = ops.pack_split_weight(_pack_conv(_fold_wn(sd, f"{ru}.1", dev), F32), C)
                u["w1s"] = ops.pack_split_weight(_pack_conv(_fold_wn(sd, f"{ru}.3", dev), F32), C)
            blk["wds"] = ops.pack_split_weight(_pack_conv_strided(_fold_wn(sd, f"{b}.4", dev), blk["rate"], F32), C)
            C *= 2
        self.enc_out_ws = ops.pack_split_weight(_pack_conv(_fold_wn(sd, f"{eb}.{n + 2}", dev), F32), C)
        D = cfg.latent_dim
        for u, dw in enumerate(self.downs):
            p = f"quantizer.downsample.{u}"
            dw["wds"] = ops.pack_split_weight(_pack_conv_strided(sd[f"{p}.0.conv.weight"].to(dev).float(), dw["factor"], F32), D)
            w1, w2 = sd[f"{p}.1.pwconv1.weight"].to(dev).float(), sd[f"{p}.1.pwconv2.weight"].to(dev).float()
            dw["pw1s"], dw["pw2s"] = ops.pack_split_weight(w1, w1.shape[1]), ops.pack_split_weight(w2, w2.shape[1])

        def tr_split(tw, pm, n_layers):
            for i in range(n_layers):
                p, lw = f"{pm}.layers.{i}", tw["layers"][i]
                g = lambda k: sd[f"{p}.{k}.weight"].to(dev).float()
                lw["wqkvs"] = ops.pack_split_weight(g("attention.wqkv"), g("attention.wqkv").shape[1])
                lw["wos"] = ops.pack_split_weight(g("attention.wo"), g("attention.wo").shape[1])
                w13 = torch.cat([g("feed_forward.w1"), g("feed_forward.w3")], 0)              # un-interleaved [w1; w3]
                lw["w13s"] = ops.pack_split_weight(w13, w13.shape[1])
                lw["w2s"] = ops.pack_split_weight(g("feed_forward.w2"), g("feed_forward.w2").shape[1])
        if self.enc_tr is not None:
            tr_split(self.enc_tr, f"{eb}.{n}.block.5", cfg.enc_tr_layers)
        tr_split(self.pre_tr, "quantizer.pre_module", cfg.tr_layers)
        for tw in (self.enc_tr, self.pre_tr):        # the plain bf16 operands serve the decode half's post_module only
            for lw in (tw["layers"] if tw is not None else []):
                for k in ("wqkv", "wo", "w13", "w2"):
                    del lw[k]

    @property
    def device(self):
        return self._device

    @property
    def dtype(self):
        return self._dtype

    def eval(self):
        return self

    # ------------------------------------------------------------------------------------------
    def _transformer(self, x: torch.Tensor, B: int, T: int, tw: Optional[dict] = None, *, heads: Optional[int] = None,
                     ffn: Optional[int] = None, window: Optional[int] = None, out_f32: bool = False) -> torch.Tensor:
        """WindowLimitedTransformer (autoencoder.py:744-802): x fp32 [B*T, D] residual stream (updated in place) ->
        final RMSNorm as bf16 [B*T, D] (a GEMM operand) or fp32 (the quantizer input)."""
        cfg, dev = self.cfg, self._device
        tw = self.post_tr if tw is None else tw
        D = x.shape[1]
        H, Dh = heads or cfg.tr_heads, cfg.tr_head_dim
        ffn, window = ffn or cfg.tr_ffn, window or cfg.tr_window
        M = B * T
        assert T <= tw["rope"][0].shape[0], "sequence longer than the transformer's RoPE table (block_size)"
        xn = torch.empty((M, D), device=dev, dtype=BF16)
        qkv = torch.empty((3, M, H * Dh), device=dev, dtype=BF16)
        ao = torch.empty((M, H * Dh), device=dev, dtype=BF16)
        hmid = torch.empty((M, ffn), device=dev, dtype=BF16)
        for lw in tw["layers"]:
            ops.rmsnorm_mod(x, lw["an"], None, xn, cfg.tr_eps)
            ops.gemm_sections(xn, lw["wqkv"], [(L.SEC_ROPE, qkv[0], None), (L.SEC_ROPE, qkv[1], None),
                                               (L.SEC_PLAIN, qkv[2], None)],
                              head_dim=Dh, rope_heads=H, rope=tw["rope"], rows_per_seq=T, eps=cfg.tr_eps)
            seg = ops.Segment(qkv[1].view(B, T, H, Dh), qkv[2].view(B, T, H, Dh), causal=True, window=window)
            ops.attention(qkv[0].view(B, T, H, Dh), [seg], ao)
            ops.gemm_generic(ao, lw["wo"], gate=lw["ga"], resid=x, out_f32=x)
            ops.rmsnorm_mod(x, lw["fn"], None, xn, cfg.tr_eps)
            ops.gemm_swiglu(xn, lw["w13"], hmid)
            ops.gemm_generic(hmid, lw["w2"], gate=lw["gf"], resid=x, out_f32=x)
        if out_f32:
            y = torch.empty((M, D), device=dev, dtype=F32)
            ops.rmsnorm_f32(x, tw["norm"], y, cfg.tr_eps)
            return y
        ops.rmsnorm_mod(x, tw["norm"], None, xn, cfg.tr_eps)
        return xn

    def _convnext(self, x: torch.Tensor, cw: dict, B: int, T: int, *, out_bf16=None, out_f32=None) -> None:
        """ConvNeXtBlock (autoencoder.py:340-373): x fp32 [B*T, D] -> x + gamma * pw2(gelu(pw1(LN(dwconv(x)))))."""
        dev = self._device
        D = x.shape[1]
        y = torch.empty((B * T, D), device=dev, dtype=BF16)
        ops.dwconv_layernorm(x.view(B, T, D), cw["dw"], cw["dwb"], cw["lnw"], cw["lnb"], 1e-6, y.view(B, T, D))
        hmid = torch.empty((B * T, cw["pw1"].shape[0]), device=dev, dtype=BF16)
        ops.gemm_generic(y, cw["pw1"], bias=cw["pb1"], act=L.ACT_GELU, out_bf16=hmid)
        ops.gemm_generic(hmid, cw["pw2"], bias=cw["pb2"], gate=cw["gamma"], resid=x, out_bf16=out_bf16, out_f32=out_f32)

    # ------------------------------------------------------------------------------------------
    # encode half
    # ------------------------------------------------------------------------------------------
    # ------------------------------------------------------------------------------------------
    # fp32-class encode (the path encode_zq runs): every contraction as a three-product split-bf16 GEMM over fp32
    # activations (csrc/precise.cu) so that the RVQ codes -- an integer decision -- equal the reference's fp32 run
    # ------------------------------------------------------------------------------------------
    def _transformer_split(self, x: torch.Tensor, B: int, T: int, tw: dict, *, heads: int, ffn: int, window: int) -> torch.Tensor:
        """WindowLimitedTransformer (autoencoder.py:744-802) on an fp32 residual stream [B*T, D] (in place) -> final
        RMSNorm fp32 [B*T, D]."""
        cfg, dev = self.cfg, self._device
        M, D = x.shape
        Dh = cfg.tr_head_dim
        assert heads * Dh == D and T <= tw["rope"][0].shape[0]
        xn = torch.empty((M, D), device=dev, dtype=F32)
        a = torch.empty((M, 3 * D), device=dev, dtype=BF16)
        qkv = torch.empty((M, 3 * D), device=dev, dtype=F32)
        ao = torch.empty((M, D), device=dev, dtype=F32)
        h13 = torch.empty((M, 2 * ffn), device=dev, dtype=F32)
        ah = torch.empty((M, 3 * ffn), device=dev, dtype=BF16)
        for lw in tw["layers"]:
            ops.rmsnorm_f32(x, lw["an"], xn, cfg.tr_eps)
            ops.split_bf16x3(xn, a)
            ops.gemm_generic(a, lw["wqkvs"], out_f32=qkv)
            ops.rope_f32(qkv[:, :D], heads, Dh, tw["rope"], T)
            ops.rope_f32(qkv[:, D:2 * D], heads, Dh, tw["rope"], T)
            ops.attention_window_f32(qkv[:, :D], qkv[:, D:2 * D], qkv[:, 2 * D:], ao, batch=B, heads=heads, head_dim=Dh,
                                     window=window)
            ops.split_bf16x3(ao, a)
            ops.gemm_generic(a, lw["wos"], gate=lw["ga"], resid=x, out_f32=x)
            ops.rmsnorm_f32(x, lw["fn"], xn, cfg.tr_eps)
            ops.split_bf16x3(xn, a)
            ops.gemm_generic(a, lw["w13s"], out_f32=h13)
            ops.split_bf16x3(h13, ah, mode=ops.SPLIT_SWIGLU)
            ops.gemm_generic(ah, lw["w2s"], gate=lw["gf"], resid=x, out_f32=x)
        ops.rmsnorm_f32(x, tw["norm"], xn, cfg.tr_eps)
        return xn

    def _encoder_split(self, audio: torch.Tensor):
        """Encoder.forward (autoencoder.py:903-929): audio fp32 [B, T] -> (z fp32 [B*T', latent], T')."""
        cfg, dev = self.cfg, self._device
        B, T = audio.shape
        C = self.enc_in_w.shape[0]
        x = torch.empty((B, T, C), device=dev, dtype=F32)
        scratch = torch.empty((B, T, C), device=dev, dtype=BF16)
        ops.conv_in_snake(audio, self.enc_in_w, self.enc_in_b, self.enc_blocks[0]["units"][0]["a1"], x, scratch)
        del scratch
        x = x.view(B * T, C)
        for blk in self.enc_blocks:
            r = blk["rate"]
            t = torch.empty_like(x)
            a = torch.empty((B * T, 3 * C), device=dev, dtype=BF16)
            for u in blk["units"]:                               # ResidualUnit (autoencoder.py:879-900)
                d = u["dil"]
                ops.split_bf16x3(x, a, mode=ops.SPLIT_SNAKE, alpha=u["a1"])
                ops.gemm_generic(a, u["w7s"], taps=tuple(-(6 - k) * d for k in range(7)), batch=B, bias=u["b7"], out_f32=t)
                ops.split_bf16x3(t, a, mode=ops.SPLIT_SNAKE, alpha=u["a2"])
                ops.gemm_generic(a, u["w1s"], batch=B, bias=u["b1"], resid=x, out_f32=x)
            ops.split_bf16x3(x, a, mode=ops.SPLIT_SNAKE, alpha=blk["alpha_out"])
            del t, x
            T, C = T // r, 2 * C
            x = torch.empty((B * T, C), device=dev, dtype=F32)
            ops.gemm_generic(a.view(B * T, r * 3 * (C // 2)), blk["wds"], taps=(-1, 0), batch=B, bias=blk["bd"], out_f32=x)
            del a
        if self.enc_tr is not None:
            x = self._transformer_split(x, B, T, self.enc_tr, heads=cfg.enc_tr_heads, ffn=3 * C, window=cfg.enc_tr_window)
        a = ops.split_bf16x3(x, mode=ops.SPLIT_SNAKE, alpha=self.enc_alpha_out)
        z = torch.empty((B * T, cfg.latent_dim), device=dev, dtype=F32)
        ops.gemm_generic(a, self.enc_out_ws, taps=(-2, -1, 0), batch=B, bias=self.enc_out_b, out_f32=z)
        return z, T

    def _downsample_split(self, z: torch.Tensor, B: int, T: int):
        """quantizer.downsample (autoencoder.py:416-425): z fp32 [B*T, D] -> fp32 [B*T/prod(factors), D]."""
        dev = self._device
        D = z.shape[1]
        x = z
        for dw in self.downs:
            fac = dw["factor"]
            a = ops.split_bf16x3(x)
            T //= fac
            x = torch.empty((B * T, D), device=dev, dtype=F32)
            ops.gemm_generic(a.view(B * T, fac * 3 * D), dw["wds"], batch=B, bias=dw["bd"], out_f32=x)   # k == stride: one tap
            y = torch.empty((B * T, D), device=dev, dtype=F32)                                      # ConvNeXtBlock :340-373
            ops.dwconv_layernorm_f32(x.view(B, T, D), dw["dw"], dw["dwb"], dw["lnw"], dw["lnb"], 1e-6, y.view(B, T, D))
            hmid = torch.empty((B * T, dw["pw1s"].shape[0]), device=dev, dtype=F32)
            ops.gemm_generic(ops.split_bf16x3(y), dw["pw1s"], bias=dw["pb1"], act=L.ACT_GELU, out_f32=hmid)
            ops.gemm_generic(ops.split_bf16x3(hmid), dw["pw2s"], bias=dw["pb2"], gate=dw["gamma"], resid=x, out_f32=x)
        return x, T

    def _rvq(self, z: torch.Tensor):
        """Semantic VQ + residual VQ stack in fp32 (autoencoder.py:457-468, 185-222, 129-157) and from_codes
        (:224-233): z fp32 [M, D] -> (codes int32 [1 + n_codebooks, M], z_q fp32 [M, D])."""
        dev = self._device
        M, D = z.shape
        residual = z.clone()
        nq = len(self.quantizers)
        codes = torch.empty((nq, M), device=dev, dtype=torch.int32)
        z_q = torch.empty((M, D), device=dev, dtype=F32)
        dim = self.quantizers[0]["cb"].shape[1]
        e = torch.empty((M, dim), device=dev, dtype=F32)
        code = torch.empty((M, dim), device=dev, dtype=F32)
        for i, q in enumerate(self.quantizers):
            ops.sgemm(residual, q["w_in"], bias=q["b_in"], out=e)
            ops.vq_nearest(e, q["cbn"], q["cb"], codes[i], code)
            ops.sgemm(code, q["w_out"], bias=q["b_out"], addend=(z_q if i > 0 else None), out=z_q)
            if i + 1 < nq:      # residual -= out_proj(code): exact negation of weights and bias, then add
                ops.sgemm(code, q["w_out_neg"], bias=q["b_out_neg"], addend=residual, out=residual)
        return codes, z_q

    @torch.inference_mode()
    def encode_stages(self, audio_data: torch.Tensor) -> dict:
        """encode_zq with the stage tensors the parity tests pin (reference layouts: channels-first)."""
        if not self.has_encoder:
            raise NotImplementedError("this DAC was loaded from a decode-only checkpoint (no encoder.* / quantizer "
                                      "tensors): encode_zq needs the full S1-DAC state dict")
        if not audio_data.is_cuda:
            audio_data = audio_data.to(self._device)
        a = audio_data.unsqueeze(1) if audio_data.ndim == 2 else audio_data
        assert a.ndim == 3 and a.shape[1] == 1
        cfg, dev = self.cfg, self._device
        B, _, n = a.shape
        hop = cfg.hop
        T = -(-n // hop) * hop                                     # right pad to whole frames (autoencoder.py:1091-1092)
        audio = torch.zeros((B, T), device=dev, dtype=F32)
        audio[:, :n] = a[:, 0].to(dev, F32)
        z_enc, Te = self._encoder_split(audio)
        x, Tq = self._downsample_split(z_enc, B, Te)
        z_pre = self._transformer_split(x, B, Tq, self.pre_tr, heads=cfg.tr_heads, ffn=cfg.tr_ffn, window=cfg.tr_window)
        codes, z_q = self._rvq(z_pre)
        cf = lambda t, Tn: t.view(B, Tn, -1).transpose(1, 2)
        return dict(z_enc=cf(z_enc, Te), z_pre=cf(z_pre, Tq), z_q=cf(z_q, Tq), z_q_rows=z_q,
                    codes=codes.view(-1, B, Tq).permute(1, 0, 2).long())

    def encode_zq(self, audio_data: torch.Tensor) -> torch.Tensor:
        """Reference signature (autoencoder.py:1116-1126): audio [B, 1, n] -> z_q [B, latent_dim, ceil(n / hop)]."""
        return self.encode_stages(audio_data)["z_q"].to(self._dtype)

    @torch.inference_mode()
    def encode_latent(self, audio: torch.Tensor, pca_components: torch.Tensor, pca_mean: torch.Tensor,
                      latent_scale: float) -> torch.Tensor:
        """ae_encode (inference.py:223-229) fused: the PCA projection reads z_q time-major (no transposes):
        ((z_q - mean) @ components^T) * scale = (z_q @ components^T - components @ mean) * scale, fp32 [B, T, 80]."""
        dev = self._device
        st = self.encode_stages(audio)
        comp = pca_components.to(dev, F32).contiguous()
        nb = -(comp @ pca_mean.to(dev, F32)).contiguous()
        B, _, Tq = st["z_q"].shape
        out = torch.empty((B * Tq, comp.shape[0]), device=dev, dtype=F32)
        ops.sgemm(st["z_q_rows"], comp, bias=nb, out=out, out_scale=float(latent_scale))
        return out.view(B, Tq, -1)

    def _upsample(self, a: torch.Tensor, B: int, T: int):
        """quantizer.upsample (autoencoder.py:427-435): a bf16 [B*T, D] -> bf16 [B*T*prod(factors), D]."""
        cfg, dev = self.cfg, self._device
        D = cfg.latent_dim
        for up in self.ups:
            fac = up["factor"]
            x = torch.empty((B * T, fac * D), device=dev, dtype=F32)
            ops.gemm_generic(a, up["wt"], bias=up["bt"], out_f32=x)               # k == stride: one tap, no trim
            T = T * fac
            x = x.view(B * T, D)
            y = torch.empty((B * T, D), device=dev, dtype=BF16)
            ops.dwconv_layernorm(x.view(B, T, D), up["dw"], up["dwb"], up["lnw"], up["lnb"], 1e-6, y.view(B, T, D))
            hmid = torch.empty((B * T, up["pw1"].shape[0]), device=dev, dtype=BF16)
            ops.gemm_generic(y, up["pw1"], bias=up["pb1"], act=L.ACT_GELU, out_bf16=hmid)
            a = torch.empty((B * T, D), device=dev, dtype=BF16)
            ops.gemm_generic(hmid, up["pw2"], bias=up["pb2"], gate=up["gamma"], resid=x, out_bf16=a)
        return a, T

    def _wave_decoder(self, a: torch.Tensor, B: int, T: int) -> torch.Tensor:
        """Decoder (autoencoder.py:971-998): a bf16 [B*T, latent_dim] -> waveform fp32 [B, T*prod(rates)]."""
        dev = self._device
        C = self.conv0_w.shape[0]
        s = torch.empty((B * T, C), device=dev, dtype=BF16)
        ops.gemm_generic(a, self.conv0_w, taps=tuple(-(6 - k) for k in range(7)), batch=B, bias=self.conv0_b,
                         out2_bf16=s, snake_alpha=self.blocks[0]["alpha_in"])
        for bi, blk in enumerate(self.blocks):
            r = blk["rate"]
            Cout = blk["wt"].shape[0] // r
            units = blk["units"]
            x = torch.empty((B * T, r * Cout), device=dev, dtype=F32)
            s1 = torch.empty((B * T, r * Cout), device=dev, dtype=BF16)
            ops.gemm_generic(s, blk["wt"], taps=(0, -1), batch=B, bias=blk["bt"], out_f32=x, out2_bf16=s1,
                             snake_alpha=units[0]["a1"].repeat(r).contiguous())
            T = T * r
            x, s1 = x.view(B * T, Cout), s1.view(B * T, Cout)
            s2 = torch.empty_like(s1)
            for j, u in enumerate(units):
                d = u["dil"]
                ops.gemm_generic(s1, u["w7"], taps=tuple(-(6 - k) * d for k in range(7)), batch=B, bias=u["b7"],
                                 out2_bf16=s2, snake_alpha=u["a2"])
                if j + 1 < len(units):
                    nxt = units[j + 1]["a1"]
                elif bi + 1 < len(self.blocks):
                    nxt = self.blocks[bi + 1]["alpha_in"]
                else:
                    nxt = self.alpha_out
                ops.gemm_generic(s2, u["w1"], batch=B, bias=u["b1"], resid=x, out2_bf16=s1, snake_alpha=nxt,
                             out_f32=(x if j + 1 < len(units) else None))   # after the last unit only snake(x) is read
            s = s1
        wav = torch.empty((B, T), device=dev, dtype=F32)
        ops.conv_out_tanh(s.view(B, T, -1), self.w_last, self.b_last, wav)
        return wav

    @torch.inference_mode()
    def decode_zq(self, z_q: torch.Tensor) -> torch.Tensor:
        """Reference signature: z_q [B, latent_dim, T] -> waveform [B, 1, T*hop]."""
        B, D, T = z_q.shape
        x = z_q.to(self._device, F32).transpose(1, 2).contiguous()
        if x.data_ptr() == z_q.data_ptr():       # the decode clobbers x (it is post_module's residual stream)
            x = x.clone()
        x = x.view(B * T, D)
        return self._decode_rows(x, B, T).unsqueeze(1).to(self._dtype)

    def pca_operands(self, pca_components: torch.Tensor, pca_mean: torch.Tensor):
        """(components^T fp32 [1024, 80], mean fp32 [1024]) on the device, cached per PCAState."""
        dev = self._device
        key = (pca_components.data_ptr(), pca_mean.data_ptr())
        if self._pca_cache is None or self._pca_cache[0] != key:
            self._pca_cache = (key, pca_components.to(dev, F32).t().contiguous(), pca_mean.to(dev, F32).contiguous())
        return self._pca_cache[1], self._pca_cache[2]

    # ------------------------------------------------------------------------------------------
    # native whole-decode entry points (csrc/ae_decode.cu): the launch sequence of _transformer / _upsample /
    # _wave_decoder below as ONE C call over caller-lent workspace; ECHO_AE_NATIVE=0 keeps the Python sequence (A/B).
    # ------------------------------------------------------------------------------------------
    @staticmethod
    def use_native() -> bool:
        return os.environ.get("ECHO_AE_NATIVE", "1") != "0"

    def native_desc(self) -> "L.AeDecoderDesc":
        """echo_ae_decoder_desc over this object's packed weights (built once; host arrays kept alive on self)."""
        d = self.__dict__.get("_c_desc")
        if d is not None:
            return d
        cfg = self.cfg
        p = lambda t: t.data_ptr()
        tr = (L.AeTrLayer * len(self.post_tr["layers"]))()
        for i, lw in enumerate(self.post_tr["layers"]):
            tr[i].wqkv, tr[i].wo, tr[i].w13, tr[i].w2 = p(lw["wqkv"]), p(lw["wo"]), p(lw["w13"]), p(lw["w2"])
            tr[i].attn_norm, tr[i].ffn_norm, tr[i].attn_scale, tr[i].ffn_scale = p(lw["an"]), p(lw["fn"]), p(lw["ga"]), p(lw["gf"])
        ups = (L.AeUpsample * len(self.ups))()
        for i, u in enumerate(self.ups):
            ups[i].factor, ups[i].hidden = u["factor"], u["pw1"].shape[0]
            ups[i].wt, ups[i].bt, ups[i].dw, ups[i].dw_b = p(u["wt"]), p(u["bt"]), p(u["dw"]), p(u["dwb"])
            ups[i].ln_w, ups[i].ln_b, ups[i].pw1, ups[i].pb1 = p(u["lnw"]), p(u["lnb"]), p(u["pw1"]), p(u["pb1"])
            ups[i].pw2, ups[i].pb2, ups[i].gamma = p(u["pw2"]), p(u["pb2"]), p(u["gamma"])
        blocks = (L.AeBlock * len(self.blocks))()
        keep = []
        c_in = self.conv0_w.shape[0]
        for i, b in enumerate(self.blocks):
            r = b["rate"]
            co = b["wt"].shape[0] // r
            a1r = b["units"][0]["a1"].repeat(r).contiguous()
            keep.append(a1r)
            blocks[i].rate, blocks[i].c_in, blocks[i].c_out = r, c_in, co
            blocks[i].alpha_in, blocks[i].alpha1_rep, blocks[i].wt, blocks[i].bt = p(b["alpha_in"]), p(a1r), p(b["wt"]), p(b["bt"])
            assert len(b["units"]) == 3
            for j, u in enumerate(b["units"]):
                cu = blocks[i].units[j]
                cu.dilation, cu.alpha1, cu.alpha2 = u["dil"], p(u["a1"]), p(u["a2"])
                cu.w7, cu.b7, cu.w1, cu.b1 = p(u["w7"]), p(u["b7"]), p(u["w1"]), p(u["b1"])
            c_in = co
        d = L.AeDecoderDesc()
        d.d, d.tr_layers, d.tr_heads, d.tr_head_dim = cfg.latent_dim, len(tr), cfg.tr_heads, cfg.tr_head_dim
        d.tr_ffn, d.tr_window, d.tr_eps = cfg.tr_ffn, cfg.tr_window, cfg.tr_eps
        d.tr, d.tr_norm = tr, p(self.post_tr["norm"])
        d.rope_cos, d.rope_sin, d.rope_len = p(self.post_tr["rope"][0]), p(self.post_tr["rope"][1]), self.post_tr["rope"][0].shape[0]
        d.n_up, d.ups = len(ups), ups
        d.conv0_w, d.conv0_b, d.c0 = p(self.conv0_w), p(self.conv0_b), self.conv0_w.shape[0]
        d.n_blocks, d.blocks = len(blocks), blocks
        d.alpha_out, d.w_last, d.b_last = p(self.alpha_out), p(self.w_last), self.b_last
        gws = ops.gemm_workspace(self._device)
        d.gemm_impl, d.gemm_workspace, d.gemm_workspace_bytes = ops.GEMM_IMPL, gws.data_ptr(), gws.numel()
        self._c_desc, self._c_keep = d, (tr, ups, blocks, keep, gws)
        return d

    def native_workspace(self, B: int, T: int, stream: bool) -> torch.Tensor:
        """Scratch lent to the native decode, one buffer per (batch, frames, mode), kept for the next call."""
        cache = self.__dict__.setdefault("_c_ws", {})
        key = (B, T, stream)
        ws = cache.get(key)
        if ws is None:
            n = int(L.load().echo_ae_decode_workspace_bytes(C.byref(self.native_desc()), B, T, int(stream)))
            if n < 0:
                raise RuntimeError(f"echo_b200: {L.load().echo_last_error().decode()}")
            if len(cache) >= 8:                  # requests of many different lengths: do not hoard scratch
                cache.clear()
            ws = cache[key] = torch.empty(n, device=self._device, dtype=torch.uint8)
        return ws

    def _decode_rows(self, x: torch.Tensor, B: int, T: int) -> torch.Tensor:
        """x fp32 [B*T, latent_dim] (time-major z_q, clobbered) -> waveform fp32 [B, T*hop]."""
        if not self.use_native():
            a = self._transformer(x, B, T)
            a, T2 = self._upsample(a, B, T)
            return self._wave_decoder(a, B, T2)
        ws = self.native_workspace(B, T, False)
        wav = torch.empty((B, T * self.cfg.hop), device=self._device, dtype=F32)
        L.check(L.load().echo_ae_decode(C.byref(self.native_desc()), x.data_ptr(), B, T, wav.data_ptr(), ws.data_ptr(),
                                        ws.numel(), L.stream()))
        return wav

    def decode_stream(self):
        """Incremental decode_zq for one utterance (echo_tts_b200.streaming.DecodeStream)."""
        from .streaming import DecodeStream
        return DecodeStream(self)

    def stream_pool(self, n: int):
        """``n`` decode streams kept on this object across requests and reset for a new utterance each: their state buffers
        and captured step graphs are reused (a graph is keyed by frames per step and stream position)."""
        pool = self.__dict__.setdefault("_stream_pool", [])
        while len(pool) < n:
            pool.append(self.decode_stream())
        for st in pool[:n]:
            st.reset()
        return pool[:n]

    @torch.inference_mode()
    def decode_latent(self, z: torch.Tensor, pca_components: torch.Tensor, pca_mean: torch.Tensor,
                      latent_scale: float) -> torch.Tensor:
        """ae_decode (inference.py:232-235) fused: PCA inverse straight into the time-major residual stream.
        z fp32 [B, T, 80] -> waveform fp32 [B, 1, T*hop]."""
        B, T, lat = z.shape
        dev = self._device
        w, mean = self.pca_operands(pca_components, pca_mean)
        x = torch.empty((B * T, w.shape[0]), device=dev, dtype=F32)
        ops.sgemm(z.to(dev, F32).contiguous().view(B * T, lat), w, addend=mean, out=x, out_scale=1.0 / latent_scale)
        return self._decode_rows(x, B, T).unsqueeze(1)
