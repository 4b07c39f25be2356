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
