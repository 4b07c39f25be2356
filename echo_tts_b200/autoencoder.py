"""Fish S1-DAC decode engine (``DAC.decode_zq``, reference ``autoencoder.py:1128-1132``) on sm_100a kernels.

Activations are time-major ``[B, T, C]`` so that every convolution is an implicit GEMM over TMA-shifted row
windows (causal left padding = TMA out-of-bounds zero fill), weight-norm is folded once at load
(``autoencoder.py:90-94`` recomputes it every forward), Snake / bias / residual / GELU / layer-scale live in GEMM
epilogues, and the dense [T, T] window mask built on the CPU per call (``autoencoder.py:762-797``) is replaced by
in-kernel window arithmetic.  bf16 tensor-core operands, fp32 accumulation, fp32 residual streams.
"""
from __future__ import annotations

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


def _pack_conv(w: torch.Tensor) -> torch.Tensor:
    """Conv1d weight [Cout, Cin, K] -> GEMM B operand [Cout, K*Cin] (tap-major K axis), bf16."""
    return w.permute(0, 2, 1).reshape(w.shape[0], -1).to(BF16).contiguous()


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
        pm = "quantizer.post_module"
        fc = sd[f"{pm}.freqs_cis"].to(dev).float()              # bf16-rounded table (autoencoder.py:805-813)
        self.rope = (fc[..., 0].contiguous(), fc[..., 1].contiguous())
        self.tr_layers = []
        for i in range(cfg.tr_layers):
            p = f"{pm}.layers.{i}"
            self.tr_layers.append(dict(
                wqkv=sd[f"{p}.attention.wqkv.weight"].to(dev).to(BF16).contiguous(),
                wo=sd[f"{p}.attention.wo.weight"].to(dev).to(BF16).contiguous(),
                w13=ops.pack_w13(sd[f"{p}.feed_forward.w1.weight"].to(dev), sd[f"{p}.feed_forward.w3.weight"].to(dev)).to(BF16),
                w2=sd[f"{p}.feed_forward.w2.weight"].to(dev).to(BF16).contiguous(),
                an=f(f"{p}.attention_norm.weight"), fn=f(f"{p}.ffn_norm.weight"),
                ga=f(f"{p}.attention_layer_scale.gamma"), gf=f(f"{p}.ffn_layer_scale.gamma")))
        self.tr_norm = f(f"{pm}.norm.weight")
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

    @property
    def device(self):
        return self._device

    @property
    def dtype(self):
        return self._dtype

    def eval(self):
        return self

    def encode_zq(self, audio_data):
        raise NotImplementedError(
            "S1-DAC *encode* is not on the accelerated path yet (SURVEY.md section 8 row f1): pass speaker latents "
            "directly to the sampler, or encode the reference clip once with the upstream model.")

    # ------------------------------------------------------------------------------------------
    def _transformer(self, x: torch.Tensor, B: int, T: int) -> torch.Tensor:
        """post_module (WindowLimitedTransformer, autoencoder.py:744-802): x fp32 [B*T, D] -> bf16 [B*T, D]."""
        cfg, dev = self.cfg, self._device
        D, H, Dh = cfg.latent_dim, cfg.tr_heads, cfg.tr_head_dim
        M = B * T
        xn = torch.empty((M, D), device=dev, dtype=BF16)
        qkv = torch.empty((3, M, H * Dh), device=dev, dtype=BF16)
        ao = torch.empty((M, H * Dh), device=dev, dtype=BF16)
        hmid = torch.empty((M, cfg.tr_ffn), device=dev, dtype=BF16)
        for lw in self.tr_layers:
            ops.rmsnorm_mod(x, lw["an"], None, xn, cfg.tr_eps)
            ops.gemm_sections(xn, lw["wqkv"], [(L.SEC_ROPE, qkv[0], None), (L.SEC_ROPE, qkv[1], None),
                                               (L.SEC_PLAIN, qkv[2], None)],
                              head_dim=Dh, rope_heads=H, rope=self.rope, rows_per_seq=T, eps=cfg.tr_eps)
            seg = ops.Segment(qkv[1].view(B, T, H, Dh), qkv[2].view(B, T, H, Dh), causal=True, window=cfg.tr_window)
            ops.attention(qkv[0].view(B, T, H, Dh), [seg], ao)
            ops.gemm_generic(ao, lw["wo"], gate=lw["ga"], resid=x, out_f32=x)
            ops.rmsnorm_mod(x, lw["fn"], None, xn, cfg.tr_eps)
            ops.gemm_swiglu(xn, lw["w13"], hmid)
            ops.gemm_generic(hmid, lw["w2"], gate=lw["gf"], resid=x, out_f32=x)
        ops.rmsnorm_mod(x, self.tr_norm, None, xn, cfg.tr_eps)
        return xn

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
                ops.gemm_generic(s2, u["w1"], batch=B, bias=u["b1"], resid=x, out_f32=x, out2_bf16=s1, snake_alpha=nxt)
            s = s1
        wav = torch.empty((B, T), device=dev, dtype=F32)
        ops.conv_out_tanh(s.view(B, T, -1), self.w_last, self.b_last, wav)
        return wav

    @torch.inference_mode()
    def decode_zq(self, z_q: torch.Tensor) -> torch.Tensor:
        """Reference signature: z_q [B, latent_dim, T] -> waveform [B, 1, T*hop]."""
        B, D, T = z_q.shape
        x = z_q.to(self._device, F32).transpose(1, 2).contiguous().view(B * T, D)
        a = self._transformer(x, B, T)
        a, T2 = self._upsample(a, B, T)
        return self._wave_decoder(a, B, T2).unsqueeze(1).to(self._dtype)

    @torch.inference_mode()
    def decode_latent(self, z: torch.Tensor, pca_components: torch.Tensor, pca_mean: torch.Tensor,
                      latent_scale: float) -> torch.Tensor:
        """ae_decode (inference.py:232-235) fused: PCA inverse straight into the time-major residual stream.
        z fp32 [B, T, 80] -> waveform fp32 [B, 1, T*hop]."""
        B, T, lat = z.shape
        dev = self._device
        key = (pca_components.data_ptr(), pca_mean.data_ptr())
        if self._pca_cache is None or self._pca_cache[0] != key:
            self._pca_cache = (key, pca_components.to(dev, F32).t().contiguous(), pca_mean.to(dev, F32).contiguous())
        _, w, mean = self._pca_cache
        x = torch.empty((B * T, w.shape[0]), device=dev, dtype=F32)
        ops.sgemm(z.to(dev, F32).contiguous().view(B * T, lat), w, addend=mean, out=x, out_scale=1.0 / latent_scale)
        a = self._transformer(x, B, T)
        a, T2 = self._upsample(a, B, T)
        return self._wave_decoder(a, B, T2).unsqueeze(1)
