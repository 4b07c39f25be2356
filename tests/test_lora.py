"""LoRA-merged weight ingestion: merge == the oracle's restatement of LoRALinear.merge_weights, and --
where the upstream tree is present -- the reference model wrapped by its own ``lora.py`` computes what a plain model
carrying our merged weights computes (checkpoint written by the reference's own ``save_lora_checkpoint``)."""
import importlib
import sys

import pytest
import torch

from conftest import rel_l2
from echo_tts_b200 import schema as S
from echo_tts_b200.lora import DEFAULT_TARGETS, matches_pattern, merge_lora_state_dict
from oracle import echo_oracle as O
from oracle import ref_import


def _random_adapters(sd, targets, rank, seed=3):
    g = torch.Generator().manual_seed(seed)
    lora = {}
    for k, w in sd.items():
        name = k[: -len(".weight")] if k.endswith(".weight") else None
        if name and w.dim() == 2 and matches_pattern(name, targets):
            lora[f"{name}.lora_A"] = torch.randn(rank, w.shape[1], generator=g) * 0.05
            lora[f"{name}.lora_B"] = torch.randn(w.shape[0], rank, generator=g) * 0.05
    return lora


def test_pattern_semantics():
    assert matches_pattern("blocks.3.attention.wq", DEFAULT_TARGETS)
    assert not matches_pattern("blocks.3.attention.wk_speaker", DEFAULT_TARGETS)      # speaker path preserved
    assert not matches_pattern("text_encoder.blocks.3.attention.wq", DEFAULT_TARGETS)  # '*' is ONE component, anchored
    assert not matches_pattern("blocks.3.attention.wq.weight", DEFAULT_TARGETS)
    assert matches_pattern("a.b", ["a.*"]) and not matches_pattern("a.b.c", ["a.*"]) and not matches_pattern("axb", ["a.b"])


def test_merge_equals_oracle_and_selection_rules():
    cfg = S.ECHO_TINY
    sd = S.synth_state_dict(S.dit_schema(cfg))
    rank, alpha, strength = 4, 6.0, 0.75
    lora = _random_adapters(sd, DEFAULT_TARGETS, rank)
    assert len(lora) == 2 * 9 * cfg.num_layers
    lora["blocks.0.attention.wk_speaker.lora_A"] = torch.randn(rank, cfg.model_size)        # outside the target list
    lora["blocks.0.attention.wk_speaker.lora_B"] = torch.randn(cfg.model_size, rank)
    lora["no.such.module.lora_A"] = torch.randn(rank, 8)
    ck = dict(lora_state_dict=lora, config=dict(rank=rank, alpha=alpha, target_modules=None))
    merged = merge_lora_state_dict(sd, ck, strength)
    assert merged.keys() == sd.keys()
    changed = 0
    for k, w in sd.items():
        name = k[: -len(".weight")] if k.endswith(".weight") else ""
        if f"{name}.lora_A" in lora and matches_pattern(name, DEFAULT_TARGETS):
            want = O.lora_merged_weight(w, lora[f"{name}.lora_A"], lora[f"{name}.lora_B"], alpha * strength, rank)
            assert torch.allclose(merged[k], want, rtol=1e-6, atol=1e-7)
            assert not torch.equal(merged[k], w)
            x = torch.randn(5, w.shape[1])
            assert torch.allclose(x @ merged[k].T, O.lora_linear(x, w, lora[f"{name}.lora_A"], lora[f"{name}.lora_B"],
                                                                alpha * strength, rank), rtol=1e-4, atol=1e-5)
            changed += 1
        else:
            assert merged[k] is w
    assert changed == 9 * cfg.num_layers
    # Gradio's loader passes config.get("target_modules", []): a checkpoint without the key adapts nothing
    same = merge_lora_state_dict(sd, dict(lora_state_dict=lora, config=dict(rank=rank, alpha=alpha)))
    assert all(same[k] is sd[k] for k in sd)
    with pytest.raises(ValueError):
        merge_lora_state_dict(sd, dict(lora_state_dict=lora, config=dict(rank=rank + 1, alpha=alpha, target_modules=None)))


@pytest.mark.skipif(not ref_import.available(), reason="upstream reference tree not present")
def test_merge_vs_reference_lora_live(tmp_path):
    ns = ref_import.load()
    sys.path.insert(0, ref_import.REF_DIR)
    try:
        ref_lora = importlib.import_module("lora")
    finally:
        sys.path.remove(ref_import.REF_DIR)
        sys.modules.pop("lora", None)
    cfg = S.ECHO_TINY
    sd = S.synth_state_dict(S.dit_schema(cfg))
    rank, alpha, strength = 4, 8.0, 1.5
    targets = ["blocks.*.attention.wq", "blocks.*.attention.wo", "blocks.*.mlp.w2", "blocks.*.attention.wk_text"]
    wrapped = ref_import.build_dit(ns, cfg, {k: v.clone() for k, v in sd.items()})
    wrapped, mods = ref_lora.apply_lora_to_model(wrapped, rank=rank, alpha=alpha * strength, dropout=0.0,
                                                 target_modules=targets)                 # gradio_app.py:196-206
    g = torch.Generator().manual_seed(9)
    for m in mods.values():
        m.lora_A.data = torch.randn(m.lora_A.shape, generator=g) * 0.05
        m.lora_B.data = torch.randn(m.lora_B.shape, generator=g) * 0.05
    path = str(tmp_path / "adapter.pt")
    ref_lora.save_lora_checkpoint(wrapped, path, config=dict(rank=rank, alpha=alpha, target_modules=targets))
    merged = merge_lora_state_dict(sd, torch.load(path, map_location="cpu"), strength)
    plain = ref_import.build_dit(ns, cfg, merged)
    ids, mask, _ = O.text_ids_and_mask(["adapter check", "b"], None)
    spk = torch.randn(2, 16, 80, generator=g)
    smask = torch.ones(2, 16, dtype=torch.bool)
    x = torch.randn(2, 24, 80, generator=g)
    t = torch.tensor([0.8, 0.3])
    with torch.inference_mode():
        outs = []
        for m in (wrapped.eval(), plain):
            outs.append(m(x, t, mask, smask, m.get_kv_cache_text(ids, mask), m.get_kv_cache_speaker(spk)).float())
    assert rel_l2(outs[1], outs[0]) < 1e-5
    # and the oracle on the merged weights agrees with both
    v_o = O.dit_forward(merged, cfg, x, t, mask, smask, O.kv_cache_text(merged, cfg, ids, mask),
                        O.kv_cache_speaker(merged, cfg, spk))
    assert rel_l2(v_o, outs[0]) < 5e-5


@pytest.mark.gpu
def test_lora_model_vs_oracle_on_gpu():
    from echo_tts_b200.lora import load_model_with_lora
    cfg = S.ECHO_TINY
    sd = S.synth_state_dict(S.dit_schema(cfg))
    rank = 4
    ck = dict(lora_state_dict=_random_adapters(sd, DEFAULT_TARGETS, rank), config=dict(rank=rank, alpha=8.0, target_modules=None))
    model = load_model_with_lora(sd, ck, 1.25, cfg=cfg)
    merged = merge_lora_state_dict(sd, ck, 1.25)
    g = torch.Generator().manual_seed(2)
    ids, mask, _ = O.text_ids_and_mask(["adapter on the gpu", "x"], None)
    spk = torch.randn(2, 16, 80, generator=g)
    smask = torch.ones(2, 16, dtype=torch.bool)
    x = torch.randn(2, 40, 80, generator=g)
    t = torch.tensor([0.9, 0.3])
    ref = O.dit_forward(merged, cfg, x, t, mask, smask, O.kv_cache_text(merged, cfg, ids, mask),
                        O.kv_cache_speaker(merged, cfg, spk))
    base = O.dit_forward(sd, cfg, x, t, mask, smask, O.kv_cache_text(sd, cfg, ids, mask), O.kv_cache_speaker(sd, cfg, spk))
    got = model(x.cuda(), t.cuda(), mask.cuda(), smask.cuda(), model.get_kv_cache_text(ids.cuda(), mask.cuda()),
                model.get_kv_cache_speaker(spk.cuda())).cpu()
    assert rel_l2(got, ref) < 1e-2
    assert rel_l2(base, ref) > 5e-2          # the adapters matter at this scale: the test can tell merged from base
