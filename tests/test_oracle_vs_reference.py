"""Live pin of the CPU oracle: run the UPSTREAM modules (imported read-only from /root/reference) and the oracle
restatement on the same seeded weights and inputs.  Skipped wherever the reference tree is absent (the GPU box);
the committed fixtures under tests/golden (made by oracle/gen_golden.py from the same upstream modules) carry the
pin there."""
import pytest
import torch

from conftest import rel_l2
from echo_tts_b200 import schema as S
from oracle import echo_oracle as O
from oracle import ref_import

pytestmark = pytest.mark.skipif(not ref_import.available(), reason="upstream reference tree not present")


@pytest.fixture(scope="module")
def ns():
    return ref_import.load()


@pytest.fixture(scope="module")
def tiny(ns):
    cfg = S.ECHO_TINY
    sd = S.synth_state_dict(S.dit_schema(cfg))
    return cfg, sd, ref_import.build_dit(ns, cfg, sd)


def _inputs(seed=11, B=2, S_lat=24, T_s=20):
    g = torch.Generator().manual_seed(seed)
    ids, mask, _ = O.text_ids_and_mask(["live pin of the oracle", "short"], None)
    spk = torch.randn(B, T_s, 80, generator=g)
    smask = torch.ones(B, T_s, dtype=torch.bool)
    smask[1, 12:] = False
    x = torch.randn(B, S_lat, 80, generator=g)
    t = torch.tensor([0.85, 0.35])
    return ids, mask, spk, smask, x, t


def test_tokenizer_live(ns):
    for text in ["Hello there.", "[S2] tagged — dash…", "", "x" * 800, "naïve café 日本語"]:
        for max_len, pad in ((768, False), (None, True), (16, True)):
            ids_r, mask_r = ns.inference.get_text_input_ids_and_mask([text], max_length=max_len, pad_to_max=pad)
            ids_o, mask_o, _ = O.text_ids_and_mask([text], max_len)
            assert torch.equal(ids_r.cpu(), ids_o) and torch.equal(mask_r.cpu(), mask_o)      # bit-exact


def test_kv_caches_and_forward_live(tiny):
    cfg, sd, ref = tiny
    ids, mask, spk, smask, x, t = _inputs()
    with torch.inference_mode():
        kt_r = ref.get_kv_cache_text(ids, mask)
        ks_r = ref.get_kv_cache_speaker(spk)
        v_r = ref(x, t, mask, smask, kt_r, ks_r)
    kt_o = O.kv_cache_text(sd, cfg, ids, mask)
    ks_o = O.kv_cache_speaker(sd, cfg, spk)
    for (kr, vr), (ko, vo) in zip(kt_r + ks_r, kt_o + ks_o):
        assert kr.shape == ko.shape and rel_l2(ko, kr.float()) < 1e-5 and rel_l2(vo, vr.float()) < 1e-5
    v_o = O.dit_forward(sd, cfg, x, t, mask, smask, kt_o, ks_o)
    assert v_o.shape == v_r.shape and rel_l2(v_o, v_r.float()) < 2e-5


def test_prefix_cache_live(tiny):
    cfg, sd, ref = tiny
    g = torch.Generator().manual_seed(5)
    prefix = torch.randn(2, 16, 80, generator=g)
    with torch.inference_mode():
        kl_r = ref.get_kv_cache_latent(prefix)
    kl_o = O.kv_cache_latent(sd, cfg, prefix)
    for (kr, vr), (ko, vo) in zip(kl_r, kl_o):
        assert rel_l2(ko, kr.float()) < 1e-5 and rel_l2(vo, vr.float()) < 1e-5


def test_sampler_live(ns, tiny):
    """Same initial noise: the reference draws it from torch.Generator(rng_seed) on CPU (inference.py:473-476)."""
    cfg, sd, ref = tiny
    ids, mask, spk, smask, _, _ = _inputs()
    kw = dict(num_steps=5, cfg_scale_text=2.5, cfg_scale_speaker=4.0, cfg_min_t=0.4, cfg_max_t=1.0,
              truncation_factor=0.9, rescale_k=1.2, rescale_sigma=3.0, speaker_kv_scale=1.5, speaker_kv_max_layers=1,
              speaker_kv_min_t=0.7, sequence_length=24)
    with torch.inference_mode():
        out_r = ns.inference.sample_euler_cfg_independent_guidances(ref, spk, smask, ids, mask, 3, **kw)
    out_o = O.sample_euler_cfg_independent_guidances(sd, cfg, spk, smask, ids, mask, 3, **kw)
    assert out_o.shape == out_r.shape and rel_l2(out_o, out_r.float()) < 5e-5


def test_ae_encode_live(ns):
    """S1-DAC encode (autoencoder.py:1080-1126) at AE_TINY dims: stage tensors, bit-exact codes, ragged length,
    both attention windows active (encoder transformer 512 of 608 frames, pre_module 16 of 152)."""
    cfg = S.AE_TINY
    sd = S.synth_ae_state_dict(cfg, encode=True)
    ae = ref_import.build_ae_cfg(ns, cfg, sd)
    g = torch.Generator().manual_seed(33)
    for n in (cfg.hop * 152 - 5, cfg.hop * 3, 17):
        audio = (torch.rand(1, 1, n, generator=g) * 2 - 1) * 0.7
        r = ref_import.encode_stages(ae, audio)
        o = O.encode_stages(sd, cfg, audio)
        assert torch.equal(o["codes"], r["codes"])
        for k in ("z_enc", "z_pre", "z_q"):
            assert o[k].shape == r[k].shape and rel_l2(o[k], r[k].float()) < 1e-5


def test_sample_pipeline_live(ns, tiny):
    """inference.py:346-400 end to end (tokenise -> zero speaker -> sampler -> ae_decode -> flattening-point crop) on the
    tiny DiT + tiny S1-DAC: upstream function vs the oracle's composition, same injected x_T."""
    from functools import partial
    cfg, sd, ref = tiny
    acfg = S.AE_TINY
    asd = S.synth_ae_state_dict(acfg, encode=True)
    ae = ref_import.build_ae_cfg(ns, acfg, asd)
    comp, mean, scale = O.synth_pca(latent_dim=acfg.latent_dim)
    pca = ns.inference.PCAState(pca_components=comp, pca_mean=mean, latent_scale=scale)
    kw = dict(num_steps=4, cfg_scale_text=3.0, cfg_scale_speaker=8.0, cfg_min_t=0.5, cfg_max_t=1.0, truncation_factor=0.9,
              rescale_k=None, rescale_sigma=None, speaker_kv_scale=None, speaker_kv_max_layers=None,
              speaker_kv_min_t=None, sequence_length=24)
    noise = torch.randn(1, 24, 80, generator=torch.Generator().manual_seed(5))
    for text, pad_text, pad_spk, normalize in (("Pipeline live pin.", None, None, True), ("[S1] padded: yes", 64, 8, False)):
        real_randn = torch.randn
        torch.randn = lambda *a, **k: noise.clone()
        try:
            audio_r, text_r = ns.inference.sample_pipeline(
                ref, ae, pca, partial(ns.inference.sample_euler_cfg_independent_guidances, **kw), text, None, 3,
                pad_to_max_speaker_latent_length=pad_spk, pad_to_max_text_length=pad_text, normalize_text=normalize)
        finally:
            torch.randn = real_randn
        audio_o, text_o, _ = O.sample_pipeline(sd, cfg, asd, acfg, (comp, mean, scale), kw, text, None, 3,
                                               pad_to_max_speaker_latent_length=pad_spk, pad_to_max_text_length=pad_text,
                                               normalize_text=normalize, noise=noise)
        assert text_o == text_r
        assert audio_o.shape == audio_r.shape
        assert rel_l2(audio_o, audio_r.float()) < 1e-4
