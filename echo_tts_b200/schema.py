"""State-dict schema + deterministic synthetic weights.

The engine accepts the *reference* ``state_dict`` key names (EchoDiT: reference
``model.py:472-560``; DAC decode path: reference ``autoencoder.py:1144-1192``) so
that HF checkpoints and parity tests can share one dict.  There is no network in
the build/GPU boxes, so weights are synthetic: every tensor is drawn from its own
``torch.Generator`` seeded by ``crc32(key) ^ seed`` -- independent of module
construction order, identical on every machine, never stored in git.
"""
from __future__ import annotations

import re
import zlib
from dataclasses import dataclass, asdict
from typing import Dict, Tuple

import torch


@dataclass(frozen=True)
class DiTConfig:
    """Mirror of the kwargs hard-coded at reference ``inference.py:23-42``."""
    latent_size: int = 80
    model_size: int = 2048
    num_layers: int = 24
    num_heads: int = 16
    intermediate_size: int = 5888
    norm_eps: float = 1e-5
    text_vocab_size: int = 256
    text_model_size: int = 1280
    text_num_layers: int = 14
    text_num_heads: int = 10
    text_intermediate_size: int = 3328
    speaker_patch_size: int = 4
    speaker_model_size: int = 1280
    speaker_num_layers: int = 14
    speaker_num_heads: int = 10
    speaker_intermediate_size: int = 3328
    timestep_embed_size: int = 512
    adaln_rank: int = 256

    def kwargs(self) -> dict:
        return asdict(self)

    @property
    def head_dim(self) -> int:
        return self.model_size // self.num_heads


ECHO_BASE = DiTConfig()

# Reduced architecture used by fast parity tests (same structure, head_dim 128).
ECHO_TINY = DiTConfig(
    model_size=256, num_layers=2, num_heads=2, intermediate_size=384,
    text_model_size=256, text_num_layers=2, text_num_heads=2, text_intermediate_size=384,
    speaker_model_size=256, speaker_num_layers=2, speaker_num_heads=2,
    speaker_intermediate_size=384, timestep_embed_size=128, adaln_rank=64,
)


@dataclass(frozen=True)
class AEConfig:
    """Decode-path dimensions of ``build_ae()`` (reference ``autoencoder.py:1144-1192``)."""
    latent_dim: int = 1024
    tr_layers: int = 8
    tr_heads: int = 16
    tr_head_dim: int = 64
    tr_ffn: int = 3072
    tr_window: int = 128
    tr_block_size: int = 4096
    tr_eps: float = 1e-5
    up_factors: Tuple[int, ...] = (2, 2)
    convnext_ratio: int = 4
    decoder_dim: int = 1536
    decoder_rates: Tuple[int, ...] = (8, 8, 4, 2)
    # encode half (Encoder autoencoder.py:903-929, DownsampleResidualVectorQuantize :376-447, build_ae :1144-1192)
    encoder_dim: int = 64
    encoder_rates: Tuple[int, ...] = (2, 4, 8, 8)
    enc_tr_layers: int = 4                  # WindowLimitedTransformer after the last encoder block only
    enc_tr_window: int = 512                # fixed by EncoderBlock (autoencoder.py:858), not a build_ae() parameter
    enc_tr_block_size: int = 16384
    n_codebooks: int = 9
    codebook_size: int = 1024
    codebook_dim: int = 8
    semantic_codebook_size: int = 4096

    @property
    def enc_out_dim(self) -> int:
        return self.encoder_dim * 2 ** len(self.encoder_rates)

    @property
    def enc_tr_heads(self) -> int:
        return self.enc_out_dim // 64

    @property
    def hop(self) -> int:
        h = 1
        for f in self.up_factors:
            h *= f
        for r in self.decoder_rates:
            h *= r
        return h


AE_BASE = AEConfig()
# Small AE for fast oracle-vs-kernel tests (not constructible by the reference's build_ae()).
AE_TINY = AEConfig(latent_dim=128, tr_layers=2, tr_heads=2, tr_head_dim=64, tr_ffn=256,
                   tr_window=16, tr_block_size=512, decoder_dim=256, decoder_rates=(4, 2, 2, 2),
                   encoder_dim=16, encoder_rates=(2, 2, 2, 4), enc_tr_layers=2, enc_tr_block_size=1024,
                   n_codebooks=3, codebook_size=32, semantic_codebook_size=64)


# --------------------------------------------------------------------------------------
# key -> shape
# --------------------------------------------------------------------------------------

def _encoder_block_keys(prefix: str, d: int, h: int, f: int) -> Dict[str, Tuple[int, ...]]:
    hd = d // h
    out = {}
    for w in ("wq", "wk", "wv", "wo", "gate"):
        out[f"{prefix}.attention.{w}.weight"] = (d, d)
    out[f"{prefix}.attention.q_norm.weight"] = (h, hd)
    out[f"{prefix}.attention.k_norm.weight"] = (h, hd)
    out[f"{prefix}.mlp.w1.weight"] = (f, d)
    out[f"{prefix}.mlp.w3.weight"] = (f, d)
    out[f"{prefix}.mlp.w2.weight"] = (d, f)
    out[f"{prefix}.attention_norm.weight"] = (d,)
    out[f"{prefix}.mlp_norm.weight"] = (d,)
    return out


def dit_schema(cfg: DiTConfig = ECHO_BASE, blockwise: bool = True) -> Dict[str, Tuple[int, ...]]:
    """All EchoDiT parameter names with shapes (reference ``model.py:472-560``)."""
    s: Dict[str, Tuple[int, ...]] = {}
    d, dt, ds = cfg.model_size, cfg.text_model_size, cfg.speaker_model_size
    s["text_encoder.text_embedding.weight"] = (cfg.text_vocab_size, dt)
    for i in range(cfg.text_num_layers):
        s.update(_encoder_block_keys(f"text_encoder.blocks.{i}", dt, cfg.text_num_heads,
                                     cfg.text_intermediate_size))
    encs = ["speaker_encoder"] + (["latent_encoder"] if blockwise else [])
    for enc in encs:
        s[f"{enc}.in_proj.weight"] = (ds, cfg.latent_size * cfg.speaker_patch_size)
        s[f"{enc}.in_proj.bias"] = (ds,)
        for i in range(cfg.speaker_num_layers):
            s.update(_encoder_block_keys(f"{enc}.blocks.{i}", ds, cfg.speaker_num_heads,
                                         cfg.speaker_intermediate_size))
    s["text_norm.weight"] = (dt,)
    s["speaker_norm.weight"] = (ds,)
    if blockwise:
        s["latent_norm.weight"] = (ds,)
    s["cond_module.0.weight"] = (d, cfg.timestep_embed_size)
    s["cond_module.2.weight"] = (d, d)
    s["cond_module.4.weight"] = (3 * d, d)
    s["in_proj.weight"] = (d, cfg.latent_size)
    s["in_proj.bias"] = (d,)
    for i in range(cfg.num_layers):
        p = f"blocks.{i}"
        for w in ("wq", "wk", "wv", "gate", "wo"):
            s[f"{p}.attention.{w}.weight"] = (d, d)
        s[f"{p}.attention.wk_text.weight"] = (d, dt)
        s[f"{p}.attention.wv_text.weight"] = (d, dt)
        s[f"{p}.attention.wk_speaker.weight"] = (d, ds)
        s[f"{p}.attention.wv_speaker.weight"] = (d, ds)
        if blockwise:
            s[f"{p}.attention.wk_latent.weight"] = (d, ds)
            s[f"{p}.attention.wv_latent.weight"] = (d, ds)
        s[f"{p}.attention.q_norm.weight"] = (cfg.num_heads, cfg.head_dim)
        s[f"{p}.attention.k_norm.weight"] = (cfg.num_heads, cfg.head_dim)
        s[f"{p}.mlp.w1.weight"] = (cfg.intermediate_size, d)
        s[f"{p}.mlp.w3.weight"] = (cfg.intermediate_size, d)
        s[f"{p}.mlp.w2.weight"] = (d, cfg.intermediate_size)
        for ada in ("attention_adaln", "mlp_adaln"):
            for part in ("shift", "scale", "gate"):
                s[f"{p}.{ada}.{part}_down.weight"] = (cfg.adaln_rank, d)
                s[f"{p}.{ada}.{part}_up.weight"] = (d, cfg.adaln_rank)
                s[f"{p}.{ada}.{part}_up.bias"] = (d,)
    s["out_norm.weight"] = (d,)
    s["out_proj.weight"] = (cfg.latent_size, d)
    s["out_proj.bias"] = (cfg.latent_size,)
    return s


def ae_decode_schema(cfg: AEConfig = AE_BASE) -> Dict[str, Tuple[int, ...]]:
    """Parameters on the ``DAC.decode_zq`` path (reference ``autoencoder.py:1128-1132``)."""
    s: Dict[str, Tuple[int, ...]] = {}
    D = cfg.latent_dim
    pm = "quantizer.post_module"
    qkv = 3 * cfg.tr_heads * cfg.tr_head_dim
    for i in range(cfg.tr_layers):
        p = f"{pm}.layers.{i}"
        s[f"{p}.attention.wqkv.weight"] = (qkv, D)
        s[f"{p}.attention.wo.weight"] = (D, cfg.tr_heads * cfg.tr_head_dim)
        s[f"{p}.feed_forward.w1.weight"] = (cfg.tr_ffn, D)
        s[f"{p}.feed_forward.w3.weight"] = (cfg.tr_ffn, D)
        s[f"{p}.feed_forward.w2.weight"] = (D, cfg.tr_ffn)
        s[f"{p}.ffn_norm.weight"] = (D,)
        s[f"{p}.attention_norm.weight"] = (D,)
        s[f"{p}.attention_layer_scale.gamma"] = (D,)
        s[f"{p}.ffn_layer_scale.gamma"] = (D,)
    s[f"{pm}.norm.weight"] = (D,)
    for u, f in enumerate(cfg.up_factors):
        p = f"quantizer.upsample.{u}"
        s[f"{p}.0.conv.weight"] = (D, D, f)
        s[f"{p}.0.conv.bias"] = (D,)
        s[f"{p}.1.gamma"] = (D,)
        s[f"{p}.1.dwconv.conv.weight"] = (D, 1, 7)
        s[f"{p}.1.dwconv.conv.bias"] = (D,)
        s[f"{p}.1.norm.weight"] = (D,)
        s[f"{p}.1.norm.bias"] = (D,)
        s[f"{p}.1.pwconv1.weight"] = (cfg.convnext_ratio * D, D)
        s[f"{p}.1.pwconv1.bias"] = (cfg.convnext_ratio * D,)
        s[f"{p}.1.pwconv2.weight"] = (D, cfg.convnext_ratio * D)
        s[f"{p}.1.pwconv2.bias"] = (D,)

    C = cfg.decoder_dim
    dm = "decoder.model"

    def wnconv(prefix: str, cout: int, cin: int, k: int):
        s[f"{prefix}.conv.bias"] = (cout,)
        s[f"{prefix}.conv.parametrizations.weight.original0"] = (cout, 1, 1)
        s[f"{prefix}.conv.parametrizations.weight.original1"] = (cout, cin, k)

    def wnconvT(prefix: str, cin: int, cout: int, k: int):
        s[f"{prefix}.conv.bias"] = (cout,)
        s[f"{prefix}.conv.parametrizations.weight.original0"] = (cin, 1, 1)
        s[f"{prefix}.conv.parametrizations.weight.original1"] = (cin, cout, k)

    wnconv(f"{dm}.0", C, D, 7)
    c_in = C
    for i, r in enumerate(cfg.decoder_rates):
        c_out = c_in // 2
        b = f"{dm}.{i + 1}.block"
        s[f"{b}.0.alpha"] = (1, c_in, 1)
        wnconvT(f"{b}.1", c_in, c_out, 2 * r)
        for j in range(3):
            ru = f"{b}.{2 + j}.block"
            s[f"{ru}.0.alpha"] = (1, c_out, 1)
            wnconv(f"{ru}.1", c_out, c_out, 7)
            s[f"{ru}.2.alpha"] = (1, c_out, 1)
            wnconv(f"{ru}.3", c_out, c_out, 1)
        c_in = c_out
    n = len(cfg.decoder_rates)
    s[f"{dm}.{n + 1}.alpha"] = (1, c_in, 1)
    wnconv(f"{dm}.{n + 2}", 1, c_in, 7)
    return s


def _transformer_keys(s, pm: str, layers: int, D: int, heads: int, head_dim: int, ffn: int) -> None:
    for i in range(layers):
        p = f"{pm}.layers.{i}"
        s[f"{p}.attention.wqkv.weight"] = (3 * heads * head_dim, D)
        s[f"{p}.attention.wo.weight"] = (D, heads * head_dim)
        s[f"{p}.feed_forward.w1.weight"] = (ffn, D)
        s[f"{p}.feed_forward.w3.weight"] = (ffn, D)
        s[f"{p}.feed_forward.w2.weight"] = (D, ffn)
        s[f"{p}.ffn_norm.weight"] = (D,)
        s[f"{p}.attention_norm.weight"] = (D,)
        s[f"{p}.attention_layer_scale.gamma"] = (D,)
        s[f"{p}.ffn_layer_scale.gamma"] = (D,)
    s[f"{pm}.norm.weight"] = (D,)


def ae_encode_schema(cfg: AEConfig = AE_BASE) -> Dict[str, Tuple[int, ...]]:
    """Parameters on the ``DAC.encode_zq`` path (reference ``autoencoder.py:1080-1126``): Encoder, quantizer.downsample,
    quantizer.pre_module and the 1 + n_codebooks vector quantizers (key names as the reference's ``state_dict``)."""
    s: Dict[str, Tuple[int, ...]] = {}

    def wnconv(prefix: str, cout: int, cin: int, k: int, conv: str = ".conv"):
        s[f"{prefix}{conv}.bias"] = (cout,)
        s[f"{prefix}{conv}.parametrizations.weight.original0"] = (cout, 1, 1)
        s[f"{prefix}{conv}.parametrizations.weight.original1"] = (cout, cin, k)

    eb = "encoder.block"
    c = cfg.encoder_dim
    wnconv(f"{eb}.0", c, 1, 7)
    n = len(cfg.encoder_rates)
    for i, r in enumerate(cfg.encoder_rates):
        b = f"{eb}.{i + 1}.block"
        for j in range(3):
            ru = f"{b}.{j}.block"
            s[f"{ru}.0.alpha"] = (1, c, 1)
            wnconv(f"{ru}.1", c, c, 7)
            s[f"{ru}.2.alpha"] = (1, c, 1)
            wnconv(f"{ru}.3", c, c, 1)
        s[f"{b}.3.alpha"] = (1, c, 1)
        wnconv(f"{b}.4", 2 * c, c, 2 * r)
        c *= 2
        if i == n - 1 and cfg.enc_tr_layers > 0:
            _transformer_keys(s, f"{b}.5", cfg.enc_tr_layers, c, c // 64, 64, 3 * c)
    s[f"{eb}.{n + 1}.alpha"] = (1, c, 1)
    wnconv(f"{eb}.{n + 2}", cfg.latent_dim, c, 3)
    D = cfg.latent_dim
    for u, f in enumerate(cfg.up_factors):
        p = f"quantizer.downsample.{u}"
        s[f"{p}.0.conv.weight"] = (D, D, f)
        s[f"{p}.0.conv.bias"] = (D,)
        s[f"{p}.1.gamma"] = (D,)
        s[f"{p}.1.dwconv.conv.weight"] = (D, 1, 7)
        s[f"{p}.1.dwconv.conv.bias"] = (D,)
        s[f"{p}.1.norm.weight"] = (D,)
        s[f"{p}.1.norm.bias"] = (D,)
        s[f"{p}.1.pwconv1.weight"] = (cfg.convnext_ratio * D, D)
        s[f"{p}.1.pwconv1.bias"] = (cfg.convnext_ratio * D,)
        s[f"{p}.1.pwconv2.weight"] = (D, cfg.convnext_ratio * D)
        s[f"{p}.1.pwconv2.bias"] = (D,)
    _transformer_keys(s, "quantizer.pre_module", cfg.tr_layers, D, cfg.tr_heads, cfg.tr_head_dim, cfg.tr_ffn)
    for name, n_q, size in (("quantizer.semantic_quantizer", 1, cfg.semantic_codebook_size),
                            ("quantizer.quantizer", cfg.n_codebooks, cfg.codebook_size)):
        for q in range(n_q):
            p = f"{name}.quantizers.{q}"
            wnconv(f"{p}.in_proj", cfg.codebook_dim, D, 1, conv="")
            wnconv(f"{p}.out_proj", D, cfg.codebook_dim, 1, conv="")
            s[f"{p}.codebook.weight"] = (size, cfg.codebook_dim)
    return s


# --------------------------------------------------------------------------------------
# deterministic synthetic weights
# --------------------------------------------------------------------------------------

def _gen(key: str, seed: int) -> torch.Generator:
    g = torch.Generator(device="cpu")
    g.manual_seed((zlib.crc32(key.encode()) ^ (seed * 0x9E3779B1)) & 0x7FFFFFFF)
    return g


def _fan_in(shape: Tuple[int, ...]) -> int:
    f = 1
    for x in shape[1:]:
        f *= x
    return max(f, 1)


def synth_tensor(key: str, shape: Tuple[int, ...], seed: int, init: str = "unit") -> torch.Tensor:
    """One deterministic fp32 tensor.  The rules keep activations O(1) and make every
    parameter matter (non-unit norm weights, non-zero biases, non-trivial gates).

    ``init`` picks the scale of the dense (Linear/conv) weights: ``"unit"`` = unit-gain uniform
    U(+-sqrt(3/fan_in)) -- every sub-layer output is O(1), the hardest case for bf16 operand rounding
    (stress tests, tiny fixtures); ``"torch"`` = U(+-1/sqrt(fan_in)), the scale ``nn.Linear``'s default
    ``kaiming_uniform_(a=sqrt(5))`` gives -- i.e. what constructing the reference ``EchoDiT`` with
    random init (BASELINE.json configs[0], north_star "identical random-init weights") produces."""
    g = _gen(key, seed)
    leaf = key.rsplit(".", 1)[-1]
    if leaf == "alpha":                       # Snake1d alpha > 0
        return 0.5 + torch.rand(shape, generator=g)
    if leaf == "gamma":                       # LayerScale / ConvNeXt gamma
        return 0.2 * (0.5 + torch.rand(shape, generator=g))
    if leaf == "original0":                   # weight-norm g (positive)
        base = 0.5 + torch.rand(shape, generator=g)
        if shape[0] == 1:                     # final 1-channel conv: keep tanh in its linear range
            return 1.5 * base
        if re.search(r"block\.\d+\.block\.3\.conv", key):   # residual-unit output conv: small residual branch
            return 0.2 * base
        if re.search(r"model\.\d+\.block\.1\.conv", key):   # up-sampling transposed conv
            return 0.9 * base
        return base                           # (keeps activations O(1) so Snake's sin() is not chaotic in bf16)
    if key.endswith("norm.weight") and len(shape) <= 2:
        return 1.0 + 0.2 * (torch.rand(shape, generator=g) - 0.5)
    if leaf == "bias":
        return 0.1 * (torch.rand(shape, generator=g) - 0.5)
    if "text_embedding" in key or key.endswith("codebook.weight"):
        return torch.randn(shape, generator=g)
    bound = (3.0 / _fan_in(shape)) ** 0.5     # unit-gain uniform
    if init == "torch":
        bound = (1.0 / _fan_in(shape)) ** 0.5
    else:
        assert init == "unit", init
    if "dwconv" in key:
        bound = (3.0 / 7) ** 0.5
    return (torch.rand(shape, generator=g) * 2 - 1) * bound


def synth_state_dict(schema: Dict[str, Tuple[int, ...]], seed: int = 1234, init: str = "unit") -> Dict[str, torch.Tensor]:
    return {k: synth_tensor(k, shp, seed, init) for k, shp in schema.items()}


def ae_rope_table(cfg: AEConfig = AE_BASE, block_size: int = 0, head_dim: int = 0) -> torch.Tensor:
    """``quantizer.post_module.freqs_cis`` buffer: [block_size, head_dim/2, 2] rounded to
    **bf16** exactly as the reference builds it (``autoencoder.py:805-813``)."""
    n = head_dim or cfg.tr_head_dim
    freqs = 1.0 / (10000 ** (torch.arange(0, n, 2)[: n // 2].float() / n))
    t = torch.arange(block_size or cfg.tr_block_size)
    freqs = torch.outer(t, freqs)
    fc = torch.polar(torch.ones_like(freqs), freqs)
    return torch.stack([fc.real, fc.imag], dim=-1).to(torch.bfloat16)


def synth_ae_state_dict(cfg: AEConfig = AE_BASE, seed: int = 4321, encode: bool = False) -> Dict[str, torch.Tensor]:
    """Decode-path tensors (+ the encode path's with ``encode=True``; the decode tensors do not change)."""
    sd = synth_state_dict(ae_decode_schema(cfg), seed)
    sd["quantizer.post_module.freqs_cis"] = ae_rope_table(cfg)
    if encode:
        sd.update(synth_state_dict(ae_encode_schema(cfg), seed))
        sd["quantizer.pre_module.freqs_cis"] = ae_rope_table(cfg)
        if cfg.enc_tr_layers > 0:
            sd[f"encoder.block.{len(cfg.encoder_rates)}.block.5.freqs_cis"] = ae_rope_table(cfg, cfg.enc_tr_block_size, 64)
    return sd


def synth_pca(latent_size: int = 80, latent_dim: int = 1024, seed: int = 99):
    """Synthetic ``PCAState`` constants (``inference.py:116-135``; no network for the real file): orthonormal component
    rows, small mean, scale 0.5.  -> (components [latent_size, latent_dim], mean [latent_dim], latent_scale)."""
    g = torch.Generator().manual_seed(seed)
    q, _ = torch.linalg.qr(torch.randn(latent_dim, latent_size, generator=g))
    return q.t().contiguous(), 0.05 * torch.randn(latent_dim, generator=g), 0.5


def synth_request(B: int = 1, n_valid: int = 256, pad_to: int = 768, t_spk: int = 212, seed: int = 7):
    """Synthetic request tensors of the BASELINE.json shapes: BOS + ``n_valid - 1`` printable ASCII bytes
    padded to ``pad_to`` (``inference.py:185-217``), an N(0,1) speaker latent of ``t_spk`` frames with an all-true
    mask.  CPU generator => identical on every machine; bench.py, the parity tests and the fixture generators
    (``oracle/gen_golden_full.py``) all draw their inputs here.  -> (ids int32, text_mask bool, speaker_latent, speaker_mask)"""
    g = torch.Generator().manual_seed(seed)
    ids = torch.zeros((B, pad_to), dtype=torch.int32)
    mask = torch.zeros((B, pad_to), dtype=torch.bool)
    ids[:, 1:n_valid] = torch.randint(32, 127, (B, n_valid - 1), generator=g, dtype=torch.int32)
    mask[:, :n_valid] = True
    spk = torch.randn((B, t_spk, 80), generator=g)
    smask = torch.ones((B, t_spk), dtype=torch.bool)
    return ids, mask, spk, smask
