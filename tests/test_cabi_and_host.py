"""CPU-side checks: the C-ABI library loads and exports every symbol include/echo_b200.h declares, the ctypes
mirrors have the C struct sizes, the host-side index logic is bit-exact with the reference fixtures, and the
product path refuses to run without CUDA (no CPU fallback)."""
import json
import os
import re
import subprocess

import pytest
import torch

from conftest import GOLDEN, ROOT


def test_library_exports_every_declared_symbol():
    from echo_tts_b200 import _lib
    lib = _lib.load()
    hdr = open(os.path.join(ROOT, "include", "echo_b200.h")).read()
    declared = set(re.findall(r"\b(echo_[a-z0-9_]+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    assert declared == set(_lib.EXPORTS), declared ^ set(_lib.EXPORTS)
    for name in declared:
        assert getattr(lib, name) is not None
    assert lib.echo_version() >= 100


def test_ctypes_struct_sizes_match_c(tmp_path):
    import ctypes
    from echo_tts_b200 import _lib
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include "echo_b200.h"\nint main(){printf("%zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu\\n",'
                   'sizeof(echo_gemm_desc),sizeof(echo_sgemm_desc),sizeof(echo_attn_desc),sizeof(echo_attn_segment),'
                   'sizeof(echo_tail_desc),sizeof(echo_gemm_section),sizeof(echo_dit_layer),sizeof(echo_dit_kv),'
                   'sizeof(echo_dit_step_desc),sizeof(echo_ae_tr_layer),sizeof(echo_ae_upsample),sizeof(echo_ae_unit),'
                   'sizeof(echo_ae_block),sizeof(echo_ae_decoder_desc),sizeof(echo_enc_layer),sizeof(echo_encode_kv_desc));return 0;}')
    exe = tmp_path / "sz"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    sizes = list(map(int, subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()))
    mine = [ctypes.sizeof(c) for c in (_lib.GemmDesc, _lib.SgemmDesc, _lib.AttnDesc, _lib.AttnSegment, _lib.TailDesc,
                                       _lib.GemmSection, _lib.DitLayer, _lib.DitKV, _lib.DitStepDesc, _lib.AeTrLayer,
                                       _lib.AeUpsample, _lib.AeUnit, _lib.AeBlock, _lib.AeDecoderDesc, _lib.EncLayer, _lib.EncodeKVDesc)]
    assert sizes == mine


def test_no_cpu_fallback():
    from echo_tts_b200 import ops
    with pytest.raises(RuntimeError):
        ops.timestep_embedding(torch.zeros(2), 8) if not torch.cuda.is_available() else (_ for _ in ()).throw(RuntimeError())
    with pytest.raises(RuntimeError):
        ops.gemm_generic(torch.zeros(8, 8, dtype=torch.bfloat16), torch.zeros(8, 8, dtype=torch.bfloat16),
                         out_f32=torch.zeros(8, 8))
    if not torch.cuda.is_available():
        from echo_tts_b200.model import EchoDiT
        with pytest.raises(RuntimeError):
            EchoDiT({})


@pytest.fixture(scope="module")
def index_golden():
    with open(os.path.join(GOLDEN, "index_golden.json")) as f:
        return json.load(f)


def test_tokenizer_bit_exact(index_golden):
    from echo_tts_b200.inference import get_text_input_ids_and_mask
    for c in index_golden["tokenizer"]:
        ids, mask, nt = get_text_input_ids_and_mask([c["text"]], c["max_length"], normalize=c["normalize"],
                                                    return_normalized_text=True, pad_to_max=c["pad_to_max"])
        n = int(mask.sum())
        assert list(ids.shape) == c["shape"] and str(ids.dtype) == c["dtype"]
        assert n == c["n_valid"] and ids[0, :n].tolist() == c["ids"] and nt[0] == c["normalized"]


def test_speaker_latent_lengths_bit_exact(index_golden):
    from echo_tts_b200.inference import get_speaker_latent_and_mask, PCAState

    class FakeAE:
        dtype = torch.float32

        def encode_zq(self, audio):
            return torch.zeros(audio.shape[0], 1024, audio.shape[-1] // 2048)
    pca = PCAState(torch.zeros(80, 1024), torch.zeros(1024), 1.0)
    for c in index_golden["speaker_lengths"]:
        lat, m = get_speaker_latent_and_mask(FakeAE(), pca, torch.zeros(1, c["n_samples"]),
                                             max_speaker_latent_length=c["max_latent"], pad_to_max=c["pad_to_max"])
        assert (lat.shape[1], int(m.sum())) == (c["latent_len"], c["n_valid"]), c


def test_step_plan_matches_reference_schedule(index_golden):
    """has_cfg per step and the live speaker-KV multiplier per (step, layer), as the reference sampler produced
    them against a spy model (oracle/gen_golden.py)."""
    from echo_tts_b200.sampler import StepPlan
    for c in index_golden["schedule"]:
        plan = StepPlan("cpu", c["num_steps"], c["cfg_min_t"], c["cfg_max_t"], c["speaker_kv_scale"],
                        c["speaker_kv_max_layers"], c["speaker_kv_min_t"], 24)
        assert [3 if h else 1 for h in plan.has_cfg] == c["rows"], c
        assert torch.allclose(plan.sched[:, 0], torch.tensor(c["t"]), rtol=0, atol=0)
        for mine, ref in zip(plan.kv_mult_rows, c["kv"]):
            assert max(abs(a - b) for a, b in zip(mine, ref)) < 1e-5
    with pytest.raises(TypeError):
        from echo_tts_b200.sampler import sample_euler_cfg_independent_guidances as s

        class M:
            device = "cpu"; cfg = None
        s(M(), None, None, torch.zeros(1, 1), None, 0, 4, 1., 1., .5, 1., None, None, None, 1.5, None, None)


def test_streaming_carry_rows_host_logic():
    """_Carry (streaming.py): zero context on the first step, and a step shorter than the context keeps the tail of the
    old context in front of the new rows -- plain tensor logic, checked on CPU tensors."""
    import torch
    from echo_tts_b200.streaming import _Carry
    c = _Carry("cpu")
    lead, n = 5, 3
    hist = []
    for t in (4, 1, 2, 6):
        buf = torch.full((lead + t, 2), -1.0)
        c.load("s", buf, lead, n)
        want = ([0.0] * n + hist)[-n:]
        assert buf[lead - n:lead, 0].tolist() == want
        new = [float(len(hist) + i + 1) for i in range(t)]
        buf[lead:] = torch.tensor(new)[:, None]
        hist += new
        c.save("s", buf, lead, t, n)
        assert c.rows["s"][:, 0].tolist() == ([0.0] * n + hist)[-n:]
