"""The CPU oracle against fixtures produced by executing the upstream reference
(oracle/gen_golden.py).  Runs everywhere, including the GPU box (no /root/reference there)."""
import json
import os

import pytest
import torch

from conftest import GOLDEN, rel_l2
from echo_tts_b200 import schema as S
from oracle import echo_oracle as O

torch.set_grad_enabled(False)


@pytest.fixture(scope="module")
def index_golden():
    with open(os.path.join(GOLDEN, "index_golden.json")) as f:
        return json.load(f)


def test_tokenizer_bit_exact(index_golden):
    for c in index_golden["tokenizer"]:
        ids, mask, nt = O.text_ids_and_mask([c["text"]], c["max_length"], normalize=c["normalize"])
        assert list(ids.shape) == c["shape"] and str(ids.dtype) == c["dtype"]
        n = int(mask.sum())
        assert n == c["n_valid"] and ids[0, :n].tolist() == c["ids"] and nt[0] == c["normalized"]
        assert not ids[0, n:].any() and mask[0, :n].all()


def test_speaker_length_arithmetic(index_golden):
    for c in index_golden["speaker_lengths"]:
        _, total, valid = O.speaker_lengths(c["n_samples"], c["max_latent"], pad_to_max=c["pad_to_max"])
        assert (total, valid) == (c["latent_len"], c["n_valid"]), c
        assert c["mask_is_prefix"]


def test_flattening_point(index_golden):
    cases = torch.load(os.path.join(GOLDEN, "flatten_cases.pt"))
    for c in index_golden["flatten"]:
        assert O.find_flattening_point(cases[c["case"]]) == c["index"], c


@pytest.mark.parametrize("name,tol", [("dit_tiny_sampler.pt", 2e-5), ("dit_tiny_sampler_opts.pt", 5e-5)])
def test_sampler_matches_reference_fixture(name, tol):
    g = torch.load(os.path.join(GOLDEN, name))
    cfg = S.ECHO_TINY
    sd = S.synth_state_dict(S.dit_schema(cfg))
    trace = []
    out = O.sample_euler_cfg_independent_guidances(
        sd, cfg, g["speaker_latent"], g["speaker_mask"], g["ids"], g["text_mask"], 0,
        noise=g["noise"], trace=trace, **g["kwargs"])
    assert rel_l2(out, g["x_final"]) < tol
    assert len(trace) == len(g["v_branches"])
    for mine, ref in zip(trace, g["v_branches"]):
        assert mine["v_branches"].shape == ref.shape          # 3B rows on CFG steps, B otherwise
        assert rel_l2(mine["v_branches"], ref.float()) < tol


def test_blockwise_matches_reference_fixture():
    g = torch.load(os.path.join(GOLDEN, "dit_tiny_blockwise.pt"))
    cfg = S.ECHO_TINY
    sd = S.synth_state_dict(S.dit_schema(cfg))
    out = O.sample_blockwise_euler_cfg_independent_guidances(
        sd, cfg, g["speaker_latent"], g["speaker_mask"], g["ids"], g["text_mask"], 0,
        continuation_latent=g["continuation"], noises=g["noises"], **g["kwargs"])
    assert rel_l2(out, g["out"]) < 2e-5
    assert torch.equal(out[:, :8], g["continuation"])


@pytest.mark.slow
def test_full_arch_blockwise_matches_reference_fixture():
    """Oracle vs the upstream blockwise sampler at the full architecture (oracle/gen_golden_blockwise.py)."""
    g = torch.load(os.path.join(GOLDEN, "dit_base_blockwise.pt"))
    sd = S.synth_state_dict(S.dit_schema(S.ECHO_BASE), init=g["init"])
    out = O.sample_blockwise_euler_cfg_independent_guidances(
        sd, S.ECHO_BASE, g["speaker_latent"], g["speaker_mask"], g["ids"], g["text_mask"], 21,
        continuation_latent=g["continuation"], noises=g["noises"], **g["kwargs"])
    assert torch.equal(out[:, :16], g["continuation"])
    assert rel_l2(out, g["out"]) < 2e-5


@pytest.mark.slow
def test_full_size_block_behind_prefix_matches_reference_fixture():
    """Oracle vs the upstream blockwise sampler at BASELINE configs[4]'s size: a 160-latent block behind a 480-latent
    prefix (oracle/gen_golden_full.py c5) -- per-branch and combined velocities of both steps and the block's latent."""
    from oracle.gen_golden_full import c5_inputs
    g = torch.load(os.path.join(GOLDEN, "dit_base_c5.pt"))
    ids, mask, spk, smask, cont, noise = c5_inputs()
    assert abs(float(cont.double().sum() + noise.double().sum() + spk.double().sum()) - g["input_checksum"]) < 1e-6
    sd = S.synth_state_dict(S.dit_schema(S.ECHO_BASE), init=g["init"])
    trace = []
    out = O.sample_blockwise_euler_cfg_independent_guidances(sd, S.ECHO_BASE, spk, smask, ids, mask, 0,
                                                             continuation_latent=cont, noises=[noise], trace=trace,
                                                             **g["kwargs"])
    assert rel_l2(out[:, 480:], g["out_block"]) < 2e-5
    assert len(trace) == len(g["steps"]) == 2
    for mine, ref in zip(trace, g["steps"]):
        assert mine["has_cfg"] == ref["has_cfg"]
        assert rel_l2(mine["v_branches"], ref["v_branches"].float()) < 1e-3     # branches are stored in fp16
        assert rel_l2(mine["v"], ref["v"]) < 1e-4


@pytest.mark.slow
def test_ae_decode_matches_reference_fixture():
    g = torch.load(os.path.join(GOLDEN, "ae_base_decode.pt"))
    sd = S.synth_ae_state_dict(S.AE_BASE)
    comp, mean, scale = O.synth_pca()
    wav = O.ae_decode(sd, S.AE_BASE, comp, mean, scale, g["z"])
    assert wav.shape == g["wav"].shape == (1, 1, 12 * 2048)
    assert rel_l2(wav, g["wav"]) < 2e-4
    # causality identity: decoding a prefix equals the prefix of the decode
    wav6 = O.ae_decode(sd, S.AE_BASE, comp, mean, scale, g["z"][:, :6])
    assert rel_l2(wav6, g["wav"][..., : 6 * 2048]) < 2e-4


def _check_encode_fixture(name, cfg, with_pca=False):
    g = torch.load(os.path.join(GOLDEN, name))
    sd = S.synth_ae_state_dict(cfg, encode=True)
    st = O.encode_stages(sd, cfg, g["audio"])
    n_frames = -(-g["audio"].shape[-1] // cfg.hop)
    assert st["z_q"].shape == g["z_q"].shape == (g["audio"].shape[0], cfg.latent_dim, n_frames)
    assert st["codes"].dtype == g["codes"].dtype == torch.int64
    assert torch.equal(st["codes"], g["codes"])                      # indices: bit-exact
    for k in ("z_enc", "z_pre", "z_q"):
        assert rel_l2(st[k], g[k].float()) < 1e-3                    # fixtures are stored in fp16
    # from_codes alone (autoencoder.py:224-233) given the reference's codes
    assert rel_l2(O.zq_from_codes(sd, cfg, g["codes"]), g["z_q"].float()) < 1e-3
    if with_pca:
        comp, mean, scale = O.synth_pca()
        lat = O.ae_encode(sd, cfg, comp, mean, scale, g["audio"])
        assert lat.shape == g["latent"].shape == (1, n_frames, 80) and rel_l2(lat, g["latent"]) < 2e-5


def test_ae_encode_tiny_matches_reference_fixture():
    _check_encode_fixture("ae_encode_tiny.pt", S.AE_TINY)


@pytest.mark.slow
def test_ae_encode_base_matches_reference_fixture():
    _check_encode_fixture("ae_encode_base.pt", S.AE_BASE, with_pca=True)
