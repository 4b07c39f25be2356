"""LoRA-merged weight ingestion: the reference's Gradio path wraps target ``nn.Linear`` layers in
``LoRALinear`` (``lora.py:17-64``: y = W x + (alpha / rank) * B (A x)) after ``apply_lora_to_model``
(``lora.py:107-171``) with ``alpha`` pre-multiplied by the UI's strength slider (``gradio_app.py:196-206``).  The
packed-weight engine cannot execute a wrapped layer, so adapters are folded into the base weights *before* packing,
W' = W + (alpha * strength / rank) * B @ A -- the reference's own ``LoRALinear.merge_weights`` (``lora.py:66-82``).
Weight preparation only (once per checkpoint, plain tensor algebra on the model's device); the sampling path is
unchanged."""
from __future__ import annotations

import re
from typing import Dict, Iterable, List, Optional

import torch

# apply_lora_to_model's defaults when target_modules is None (lora.py:127-141): decoder self-attention, text
# cross-attention and the MLP; the speaker path is left alone
DEFAULT_TARGETS = (
    "blocks.*.attention.wq", "blocks.*.attention.wk", "blocks.*.attention.wv", "blocks.*.attention.wo",
    "blocks.*.attention.wk_text", "blocks.*.attention.wv_text",
    "blocks.*.mlp.w1", "blocks.*.mlp.w2", "blocks.*.mlp.w3",
)


def matches_pattern(name: str, patterns: Iterable[str]) -> bool:
    """Glob match of a dotted module name, ``*`` = exactly one name component (lora.py:85-93)."""
    for pat in patterns:
        rx = "".join("[^.]+" if ch == "*" else re.escape(ch) for ch in pat)
        if re.fullmatch(rx, name):
            return True
    return False


def merge_lora_state_dict(state_dict: Dict[str, torch.Tensor], checkpoint: dict, lora_strength: float = 1.0,
                          device=None) -> Dict[str, torch.Tensor]:
    """Base ``state_dict`` (reference key names) + a checkpoint written by the reference's ``save_lora_checkpoint``
    (``{"lora_state_dict": {"<module>.lora_A": [r, in], "<module>.lora_B": [out, r]}, "config": {...}}``,
    lora.py:196-221) -> new state dict with the adapters merged.  Same selection rules as the Gradio loader
    (gradio_app.py:183-206): rank / alpha / target_modules come from the checkpoint's config (defaults 32 / 32.0 / no
    targets; ``target_modules: None`` means the reference's default target list); adapters of modules outside the
    target list, or absent from the model, are ignored exactly as ``load_lora_checkpoint`` ignores them."""
    cfg = checkpoint.get("config", {}) or {}
    rank = int(cfg.get("rank", 32))
    alpha = float(cfg.get("alpha", 32.0)) * float(lora_strength)
    targets: Optional[List[str]] = cfg.get("target_modules", [])
    if targets is None:
        targets = list(DEFAULT_TARGETS)
    lora = checkpoint["lora_state_dict"]
    out = dict(state_dict)
    scaling = alpha / rank
    for key, A in lora.items():
        if not key.endswith(".lora_A"):
            continue
        name = key[: -len(".lora_A")]
        wkey = f"{name}.weight"
        if not matches_pattern(name, targets) or wkey not in state_dict or state_dict[wkey].dim() != 2:
            continue
        W = state_dict[wkey]
        B = lora.get(f"{name}.lora_B")
        if B is None:
            continue            # load_lora_checkpoint leaves the zero-initialised B in place: no change
        if A.shape != (rank, W.shape[1]) or B.shape != (W.shape[0], rank):
            raise ValueError(f"LoRA shapes for {name}: A {tuple(A.shape)}, B {tuple(B.shape)} do not fit weight "
                             f"{tuple(W.shape)} at rank {rank}")
        dev = W.device if device is None else torch.device(device)
        delta = (B.to(dev, torch.float32) @ A.to(dev, torch.float32)) * scaling
        out[wkey] = (W.to(dev, torch.float32) + delta).to(W.dtype)
    return out


def load_model_with_lora(state_dict: Dict[str, torch.Tensor], lora_path_or_checkpoint, lora_strength: float = 1.0, **kw):
    """``gradio_app.load_lora_model`` for the packed engine: fresh base weights + adapters -> EchoDiT.  Hot swap =
    call again with another checkpoint (the base ``state_dict`` is not modified)."""
    from .inference import load_model_from_state_dict
    ck = lora_path_or_checkpoint
    if not isinstance(ck, dict):
        ck = torch.load(ck, map_location="cpu")
    return load_model_from_state_dict(merge_lora_state_dict(state_dict, ck, lora_strength), **kw)
