"""Generate tests/golden/* by EXECUTING the upstream reference (build container only).

    python oracle/gen_golden.py [--full]      # --full also regenerates the 11 GB full-arch fixture

The reference has no tests/golden vectors of its own; these fixtures are the
pins.  Weights are never stored: they are re-synthesised from echo_tts_b200.schema (deterministic).
Texts used for tokenizer goldens are written here (not copied from the reference's presets).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import ref_import  # noqa: E402
from echo_tts_b200 import schema as S  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")

TEXTS = [
    "Hello, this is a test of the Echo TTS model.",
    "[S1] Already tagged: keep; the punctuation — mostly.",
    "(whispering) Parenthesis start… and ‘curly’ “quotes” ”here”.",
    "Line one\nline two: with colon; and semicolon",
    "S2 appears inside so no tag is added",
    "",
    "Ünïcödé — naïve café 日本語 🙂",
    "x" * 900,
]

SAMPLER_GRID = [
    dict(cfg_min_t=0.5, cfg_max_t=1.0, speaker_kv_scale=None, speaker_kv_max_layers=None, speaker_kv_min_t=None),
    dict(cfg_min_t=0.5, cfg_max_t=1.0, speaker_kv_scale=1.5, speaker_kv_max_layers=24, speaker_kv_min_t=0.9),
    dict(cfg_min_t=0.0, cfg_max_t=1.0, speaker_kv_scale=1.5, speaker_kv_max_layers=12, speaker_kv_min_t=0.5),
    dict(cfg_min_t=0.3, cfg_max_t=0.8, speaker_kv_scale=2.0, speaker_kv_max_layers=None, speaker_kv_min_t=0.25),
    dict(cfg_min_t=0.999, cfg_max_t=1.0, speaker_kv_scale=None, speaker_kv_max_layers=None, speaker_kv_min_t=None),
]


class _SpyModel:
    """Duck-typed model (surface listed in SURVEY.md section 8b) that records what the reference
    sampler asks for: rows per forward (3B => CFG step) and the live speaker-KV multiplier per layer."""
    device = torch.device("cpu")
    dtype = torch.float32

    def __init__(self, layers=24):
        self.layers = layers
        self.calls = []

    def get_kv_cache_text(self, ids, mask):
        return [(torch.ones(ids.shape[0], 1, 1, 1), torch.ones(ids.shape[0], 1, 1, 1)) for _ in range(self.layers)]

    def get_kv_cache_speaker(self, lat):
        return [(torch.ones(lat.shape[0], 1, 1, 1), torch.ones(lat.shape[0], 1, 1, 1)) for _ in range(self.layers)]

    def __call__(self, x, t, text_mask, speaker_mask, kv_cache_text, kv_cache_speaker, **kw):
        self.calls.append(dict(rows=int(x.shape[0]), t=float(t[0]),
                               kv=[round(float(k.flatten()[0]), 6) for k, _ in kv_cache_speaker]))
        return torch.zeros_like(x, dtype=torch.float32)


def index_goldens(ns) -> dict:
    inf = ns.inference
    out = {"tokenizer": [], "speaker_lengths": [], "schedule": [], "flatten": []}
    for text in TEXTS:
        for norm in (True, False):
            for max_len, pad in ((768, False), (None, True), (16, True)):
                ids, mask, nt = inf.get_text_input_ids_and_mask([text], max_length=max_len, normalize=norm,
                                                                return_normalized_text=True, pad_to_max=pad)
                out["tokenizer"].append(dict(text=text, normalize=norm, max_length=max_len, pad_to_max=pad,
                                             shape=list(ids.shape), n_valid=int(mask.sum()),
                                             ids=ids[0, : int(mask.sum())].tolist(), normalized=nt[0],
                                             dtype=str(ids.dtype)))

    class _FakeAE:                      # encode_zq stand-in: the test is about lengths/masks only
        dtype = torch.float32

        def encode_zq(self, audio):
            return torch.zeros(audio.shape[0], 1024, audio.shape[-1] // 2048)
    pca = inf.PCAState(torch.zeros(80, 1024), torch.zeros(1024), 1.0)
    for n_lat, extra in ((1, 0), (3, 5), (4, 0), (215, 700), (640, 0), (641, 1), (1300, 17), (6400, 0), (6459, 3)):
        n = n_lat * 2048 + extra
        for pad, mx in ((False, 6400), (True, 6400), (True, 2816)):
            lat, m = inf.get_speaker_latent_and_mask(_FakeAE(), pca, torch.zeros(1, n), max_speaker_latent_length=mx,
                                                     pad_to_max=pad)
            out["speaker_lengths"].append(dict(n_samples=n, pad_to_max=pad, max_latent=mx,
                                               latent_len=int(lat.shape[1]), n_valid=int(m.sum()),
                                               mask_is_prefix=bool((m[0].int().diff() <= 0).all())))
    for steps in (1, 5, 20, 40, 80):
        for g in SAMPLER_GRID:
            spy = _SpyModel()
            inf.sample_euler_cfg_independent_guidances(
                spy, torch.zeros(1, 8, 80), torch.ones(1, 8, dtype=torch.bool), torch.zeros(1, 4, dtype=torch.int32),
                torch.ones(1, 4, dtype=torch.bool), 0, num_steps=steps, cfg_scale_text=3.0, cfg_scale_speaker=8.0,
                truncation_factor=None, rescale_k=None, rescale_sigma=None, sequence_length=4, **g)
            out["schedule"].append(dict(num_steps=steps, **g, rows=[c["rows"] for c in spy.calls],
                                        t=[c["t"] for c in spy.calls], kv=[c["kv"] for c in spy.calls]))
    gen = torch.Generator().manual_seed(11)
    cases = {
        "noise": torch.randn(64, 80, generator=gen),
        "flat_from_30": torch.cat([torch.randn(30, 80, generator=gen), 0.01 * torch.randn(34, 80, generator=gen)]),
        "flat_offset_mean": torch.cat([torch.randn(10, 80, generator=gen), 0.3 + 0.01 * torch.randn(54, 80, generator=gen)]),
        "tail_partial": torch.cat([torch.randn(60, 80, generator=gen), 0.02 * torch.randn(4, 80, generator=gen)]),
        "all_flat": 0.001 * torch.randn(40, 80, generator=gen),
        "short": torch.randn(5, 80, generator=gen),
    }
    torch.save(cases, os.path.join(GOLD, "flatten_cases.pt"))
    for name, data in cases.items():
        out["flatten"].append(dict(case=name, index=int(inf.find_flattening_point(data))))
    return out


def _sampler_case(ns, cfg, seed, B, S, T_spk, steps, extra, full_trace=True, init="unit"):
    """Run the reference sampler on CPU/fp32 with injected noise, recording per-step tensors."""
    sd = S_.synth_state_dict(S_.dit_schema(cfg), init=init)
    model = ref_import.build_dit(ns, cfg, sd)
    gen = torch.Generator().manual_seed(seed)
    texts = ["Hello, this is a test of the Echo TTS model.", "Second, shorter prompt"][:B]
    ids, mask = ns.inference.get_text_input_ids_and_mask(texts, max_length=None, normalize=True)
    spk = torch.randn(B, T_spk, 80, generator=gen)
    smask = torch.ones(B, T_spk, dtype=torch.bool)
    if B > 1:
        smask[1, T_spk // 2:] = False
    noise = torch.randn(B, S, 80, generator=gen)
    kw = dict(num_steps=steps, cfg_scale_text=3.0, cfg_scale_speaker=8.0, cfg_min_t=0.5, cfg_max_t=1.0,
              truncation_factor=None, rescale_k=None, rescale_sigma=None, speaker_kv_scale=None,
              speaker_kv_max_layers=None, speaker_kv_min_t=None, sequence_length=S)
    kw.update(extra)
    trace = []
    orig_fwd = model.forward

    def spy(*a, **k):
        v = orig_fwd(*a, **k)
        trace.append(v.clone())
        return v
    model.forward = spy
    real_randn = torch.randn
    torch.randn = lambda *a, **k: noise.clone()          # inject x_T (inference.py:499)
    try:
        t0 = time.time()
        out = ns.inference.sample_euler_cfg_independent_guidances(model, spk, smask, ids, mask, seed, **kw)
        dt = time.time() - t0
    finally:
        torch.randn = real_randn
    return dict(ids=ids, text_mask=mask, speaker_latent=spk, speaker_mask=smask, noise=noise, kwargs=kw, init=init,
                v_branches=[t.half() if not full_trace else t for t in trace], x_final=out, ref_seconds=dt)


S_ = S


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--full", action="store_true")
    args = ap.parse_args()
    ns = ref_import.load()
    assert ns is not None, "reference not found"
    os.makedirs(GOLD, exist_ok=True)
    torch.manual_seed(0)

    with open(os.path.join(GOLD, "index_golden.json"), "w") as f:
        json.dump(index_goldens(ns), f, indent=0)

    # tiny architecture: plain CFG; all options on
    torch.save(_sampler_case(ns, S.ECHO_TINY, 5, 2, 32, 24, 4, {}), os.path.join(GOLD, "dit_tiny_sampler.pt"))
    torch.save(_sampler_case(ns, S.ECHO_TINY, 6, 2, 48, 16, 6,
                             dict(truncation_factor=0.8, rescale_k=1.2, rescale_sigma=3.0, speaker_kv_scale=1.5,
                                  speaker_kv_max_layers=1, speaker_kv_min_t=0.7, cfg_min_t=0.4)),
               os.path.join(GOLD, "dit_tiny_sampler_opts.pt"))

    # tiny blockwise with continuation
    cfg = S.ECHO_TINY
    sd = S.synth_state_dict(S.dit_schema(cfg))
    model = ref_import.build_dit(ns, cfg, sd)
    gen = torch.Generator().manual_seed(8)
    ids, mask = ns.inference.get_text_input_ids_and_mask(["Blockwise generation test."], max_length=None)
    spk = torch.randn(1, 16, 80, generator=gen)
    smask = torch.ones(1, 16, dtype=torch.bool)
    cont = torch.randn(1, 8, 80, generator=gen)
    noises = [torch.randn(1, b, 80, generator=gen) for b in (8, 12)]
    it = iter(noises)
    real_randn = torch.randn
    torch.randn = lambda *a, **k: next(it).clone()
    kwb = dict(block_sizes=[8, 12], num_steps=4, cfg_scale_text=3.0, cfg_scale_speaker=5.0, cfg_min_t=0.5,
               cfg_max_t=1.0, truncation_factor=0.8, rescale_k=None, rescale_sigma=None, speaker_kv_scale=None,
               speaker_kv_max_layers=None, speaker_kv_min_t=None)
    try:
        outb = ns.inference_blockwise.sample_blockwise_euler_cfg_independent_guidances(
            model, spk, smask, ids, mask, 8, continuation_latent=cont, **kwb)
    finally:
        torch.randn = real_randn
    torch.save(dict(ids=ids, text_mask=mask, speaker_latent=spk, speaker_mask=smask, continuation=cont,
                    noises=noises, kwargs=kwb, out=outb), os.path.join(GOLD, "dit_tiny_blockwise.pt"))

    # full-size AE decode of a few frames (+ causality: prefix decode equals decode prefix)
    from oracle import echo_oracle as O
    sd_ae = S.synth_ae_state_dict(S.AE_BASE)
    ae = ref_import.build_ae(ns, sd_ae)
    comp, mean, scale = O.synth_pca()
    pca = ns.inference.PCAState(pca_components=comp, pca_mean=mean, latent_scale=scale)
    z = torch.randn(1, 12, 80, generator=torch.Generator().manual_seed(3))
    wav = ns.inference.ae_decode(ae, pca, z)
    torch.save(dict(z=z, wav=wav), os.path.join(GOLD, "ae_base_decode.pt"))

    if args.full:
        # BASELINE.json config 1: full architecture, fp32 CPU, S=64, 4 steps, 212-latent speaker ref
        # (a) the contract fixture: weights at nn.Linear's default-init scale (what random-init of the reference gives)
        # (b) a stress fixture: unit-gain weights (every sub-layer O(1): worst case for bf16 operand rounding)
        for init, name in (("torch", "dit_base_c1.pt"), ("unit", "dit_base_c1_unitgain.pt")):
            case = _sampler_case(ns, S.ECHO_BASE, 7, 1, 64, 212, 4, {}, init=init)
            torch.save(case, os.path.join(GOLD, name))
            print(name, "reference sampler seconds:", case["ref_seconds"])


if __name__ == "__main__":
    main()
