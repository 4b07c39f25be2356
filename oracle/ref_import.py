"""Import the upstream reference (read-only, build container only) for golden generation and
live oracle validation.  Returns None when /root/reference is absent (e.g. on the GPU box)."""
from __future__ import annotations

import importlib
import os
import sys
import types

REF_DIR = os.environ.get("ECHO_REFERENCE_DIR", "/root/reference")


def available() -> bool:
    return os.path.isfile(os.path.join(REF_DIR, "model.py"))


def load():
    """-> namespace with .model, .autoencoder, .inference, .inference_blockwise (or None)."""
    if not available():
        return None
    # torchcodec is imported at inference.py:8 but only used by load_audio (inference.py:141-149).
    if "torchcodec" not in sys.modules:
        tc = types.ModuleType("torchcodec")
        dec = types.ModuleType("torchcodec.decoders")
        dec.AudioDecoder = object
        tc.decoders = dec
        sys.modules["torchcodec"] = tc
        sys.modules["torchcodec.decoders"] = dec
    if "torchaudio" not in sys.modules:
        try:
            importlib.import_module("torchaudio")
        except Exception:
            sys.modules["torchaudio"] = types.ModuleType("torchaudio")
    saved = {k: sys.modules.pop(k) for k in ("model", "autoencoder", "inference", "inference_blockwise")
             if k in sys.modules}
    sys.path.insert(0, REF_DIR)
    try:
        ns = types.SimpleNamespace(
            model=importlib.import_module("model"),
            autoencoder=importlib.import_module("autoencoder"),
            inference=importlib.import_module("inference"),
            inference_blockwise=importlib.import_module("inference_blockwise"),
        )
    finally:
        sys.path.remove(REF_DIR)
        for k in ("model", "autoencoder", "inference", "inference_blockwise"):
            sys.modules.pop(k, None)
        sys.modules.update(saved)
    return ns


def build_dit(ns, cfg, sd):
    """Reference EchoDiT on CPU/fp32 carrying exactly the tensors of ``sd``."""
    import torch
    with torch.device("meta"):
        m = ns.model.EchoDiT(**cfg.kwargs())
    missing = m.load_state_dict(sd, strict=False, assign=True)
    assert not missing.unexpected_keys, missing.unexpected_keys
    return m.eval()


def build_ae(ns, sd):
    """Reference DAC with the decode-path tensors of ``sd`` (encoder stays at its default init)."""
    import torch
    torch.manual_seed(0)
    ae = ns.autoencoder.build_ae()
    res = ae.load_state_dict(sd, strict=False)
    assert not res.unexpected_keys, res.unexpected_keys
    return ae.eval()


def build_ae_cfg(ns, cfg, sd):
    """Reference DAC (both halves) at the dimensions of an ``AEConfig`` -- the reference's ``build_ae()``
    (autoencoder.py:1144-1192) with the numbers taken from ``cfg`` -- carrying the tensors of ``sd``."""
    import torch
    A = ns.autoencoder
    torch.manual_seed(0)
    q_config = A.ModelArgs(block_size=cfg.tr_block_size, n_layer=cfg.tr_layers, n_head=cfg.tr_heads, dim=cfg.latent_dim,
                           intermediate_size=cfg.tr_ffn, head_dim=cfg.tr_head_dim, norm_eps=cfg.tr_eps, dropout_rate=0.1,
                           attn_dropout_rate=0.1, channels_first=True)
    make_tr = lambda: A.WindowLimitedTransformer(causal=True, window_size=cfg.tr_window, input_dim=cfg.latent_dim,
                                                 config=q_config)
    quantizer = A.DownsampleResidualVectorQuantize(
        input_dim=cfg.latent_dim, n_codebooks=cfg.n_codebooks, codebook_size=cfg.codebook_size,
        codebook_dim=cfg.codebook_dim, quantizer_dropout=0.5, downsample_factor=tuple(cfg.up_factors),
        semantic_codebook_size=cfg.semantic_codebook_size, pre_module=make_tr(), post_module=make_tr())

    def general(**kw):
        return A.ModelArgs(block_size=kw.get("block_size", cfg.enc_tr_block_size), n_layer=kw.get("n_layer", 8),
                           n_head=kw.get("n_head", 8), dim=kw.get("dim", 512),
                           intermediate_size=kw.get("intermediate_size", 1536), n_local_heads=-1, head_dim=64,
                           rope_base=10000, norm_eps=1e-5, dropout_rate=0.1, attn_dropout_rate=0.1, channels_first=True)

    n = len(cfg.encoder_rates)
    ae = A.DAC(encoder_dim=cfg.encoder_dim, encoder_rates=list(cfg.encoder_rates), latent_dim=cfg.latent_dim,
               decoder_dim=cfg.decoder_dim, decoder_rates=list(cfg.decoder_rates), quantizer=quantizer, sample_rate=44100,
               causal=True, encoder_transformer_layers=[0] * (n - 1) + [cfg.enc_tr_layers],
               decoder_transformer_layers=[0] * len(cfg.decoder_rates), transformer_general_config=general)
    res = ae.load_state_dict(sd, strict=False)
    assert not res.unexpected_keys, res.unexpected_keys
    assert all(k.endswith("causal_mask") or ".decoder." in f".{k}" and "layers" in k for k in res.missing_keys), res.missing_keys
    return ae.eval()


def encode_stages(ae, audio):
    """The reference's own encode (DAC.encode / encode_zq, autoencoder.py:1080-1126) with the intermediate tensors
    exposed: {z_enc, z_pre, codes, z_q}."""
    import math
    import torch
    import torch.nn.functional as F
    with torch.inference_mode():
        a = audio.unsqueeze(1) if audio.ndim == 2 else audio
        n = a.shape[-1]
        a = F.pad(a, (0, math.ceil(n / ae.frame_length) * ae.frame_length - n))
        z_enc = ae.encoder(a)
        z_pre = ae.quantizer.pre_module(ae.quantizer.downsample(z_enc))
        codes, _ = ae.encode(audio)
        z_q = ae.encode_zq(audio)
    return dict(z_enc=z_enc, z_pre=z_pre, codes=codes, z_q=z_q)
