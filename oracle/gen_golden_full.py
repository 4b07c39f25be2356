"""Full-size fixtures (the BENCHMARKED configurations) produced by EXECUTING the upstream reference in fp32 on CPU:

    python oracle/gen_golden_full.py [c2] [c4] [c5]      # build container only (/root/reference, ~25 GB RAM, ~20 min)

  c2  tests/golden/dit_base_c2.pt   BASELINE.json configs[1]: echo-tts-base, 1 prompt x 640 latents x 40 Euler steps, CFG
                                    3.0 / 8.0 on t in [0.5, 1], 768-padded text (256 valid), 212-latent speaker reference --
                                    the whole upstream sampler (inference.py:446-560): final latent + the sampler's x_t
                                    and per-branch / combined velocities at six selected steps (teacher forcing).
  c4  tests/golden/dit_base_c4.pt   configs[3]: B = 2, 6400-latent (5-minute) reference, Force Speaker speaker_kv_scale 1.5
                                    over 24 layers until t < 0.9; a 2-step schedule = one guided step on the SCALED cache
                                    (t = 0.999) and one un-guided step on the un-scaled cache (t = 0.4995).
  c5  tests/golden/dit_base_c5.pt   configs[4]: one 160-latent block behind a 480-latent prefix
                                    (inference_blockwise.py:14-123), cfg 3.0 / 5.0, the same 2-step schedule.

Inputs come from echo_tts_b200.schema.synth_request (CPU generators: reproducible anywhere) and are NOT stored when large;
weights are re-synthesised from the schema (seed 1234, nn.Linear-default scale) on both sides and never stored.
Branch velocities are stored in fp16 (5e-4 relative), x_t / combined velocity / final latent in fp32.
"""
from __future__ import annotations

import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import ref_import  # noqa: E402
from echo_tts_b200 import schema as S  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
C2_KEEP = (0, 10, 19, 20, 30, 39)

KW = dict(num_steps=40, cfg_scale_text=3.0, cfg_scale_speaker=8.0, cfg_min_t=0.5, cfg_max_t=1.0, truncation_factor=None,
          rescale_k=None, rescale_sigma=None, speaker_kv_scale=None, speaker_kv_max_layers=None, speaker_kv_min_t=None)


def c4_inputs():
    """configs[3] request: two prompts (different valid text lengths), 6400-latent references, the second one only
    3000 latents long (mask False and zeros behind it, inference.py:287-299)."""
    ids, mask, _, _ = S.synth_request(2, n_valid=256, seed=31)
    mask[1, 100:] = False
    ids[1, 100:] = 0
    g = torch.Generator().manual_seed(32)
    spk = torch.randn(2, 6400, 80, generator=g)
    smask = torch.ones(2, 6400, dtype=torch.bool)
    smask[1, 3000:] = False
    spk[1, 3000:] = 0.0
    noise = torch.randn(2, 640, 80, generator=g)
    return ids, mask, spk, smask, noise


def c5_inputs():
    ids, mask, spk, smask = S.synth_request(1, seed=41)
    g = torch.Generator().manual_seed(42)
    cont = torch.randn(1, 480, 80, generator=g)
    noise = torch.randn(1, 160, 80, generator=g)
    return ids, mask, spk, smask, cont, noise


def _spy(model, keep=None):
    """Record (x_in, velocities) of every model.forward the upstream sampler makes (steps in `keep` only)."""
    calls, n = [], [0]
    orig = model.forward

    def fwd(*a, **k):
        v = orig(*a, **k)
        if keep is None or n[0] in keep:
            x = k["x"] if "x" in k else a[0]
            calls.append(dict(i=n[0], x_in=x.float().clone(), v=v.float().clone()))
        n[0] += 1
        return v
    model.forward = fwd
    return calls


def _records(calls, B, cs_t, cs_s):
    out = []
    for c in calls:
        v = c["v"]
        guided = v.shape[0] == 3 * B
        rec = dict(i=c["i"], has_cfg=guided, x_in=c["x_in"][:B].clone(), v_branches=v.half())
        if guided:
            vc, vt, vs = v.chunk(3, 0)
            rec["v"] = vc + cs_t * (vc - vt) + cs_s * (vc - vs)                       # inference.py:526-530
        else:
            rec["v"] = v.clone()
        out.append(rec)
    return out


def gen_c2(ns, model):
    ids, mask, spk, smask = S.synth_request(1, seed=7)
    noise = torch.randn(1, 640, 80, generator=torch.Generator().manual_seed(11))
    calls = _spy(model, C2_KEEP)
    real = torch.randn
    torch.randn = lambda *a, **k: noise.clone()                                        # inject x_T (inference.py:499)
    try:
        t0 = time.time()
        out = ns.inference.sample_euler_cfg_independent_guidances(model, spk, smask, ids, mask, 0, sequence_length=640, **KW)
        dt = time.time() - t0
    finally:
        torch.randn = real
        del model.forward
    torch.save(dict(ids=ids, text_mask=mask, speaker_latent=spk, speaker_mask=smask, noise=noise,
                    kwargs=dict(KW, sequence_length=640), init="torch", steps=_records(calls, 1, 3.0, 8.0), x_final=out,
                    ref_seconds=dt), os.path.join(GOLD, "dit_base_c2.pt"))
    print(f"dit_base_c2.pt: upstream sampler {dt:.0f} s", flush=True)


def gen_c4(ns, model):
    ids, mask, spk, smask, noise = c4_inputs()
    kw = dict(KW, num_steps=2, speaker_kv_scale=1.5, speaker_kv_max_layers=24, speaker_kv_min_t=0.9)
    calls = _spy(model)
    real = torch.randn
    torch.randn = lambda *a, **k: noise.clone()
    try:
        t0 = time.time()
        out = ns.inference.sample_euler_cfg_independent_guidances(model, spk, smask, ids, mask, 0, sequence_length=640, **kw)
        dt = time.time() - t0
    finally:
        torch.randn = real
        del model.forward
    torch.save(dict(kwargs=dict(kw, sequence_length=640), init="torch", steps=_records(calls, 2, 3.0, 8.0), x_final=out,
                    input_checksum=float(spk.double().sum() + noise.double().sum() + ids.sum()), ref_seconds=dt),
               os.path.join(GOLD, "dit_base_c4.pt"))
    print(f"dit_base_c4.pt: upstream sampler {dt:.0f} s", flush=True)


def gen_c5(ns, model):
    ids, mask, spk, smask, cont, noise = c5_inputs()
    kw = dict(KW, num_steps=2, cfg_scale_speaker=5.0, block_sizes=[160])
    calls = _spy(model)
    real = torch.randn
    torch.randn = lambda *a, **k: noise.clone()
    try:
        t0 = time.time()
        out = ns.inference_blockwise.sample_blockwise_euler_cfg_independent_guidances(
            model, spk, smask, ids, mask, 0, continuation_latent=cont, **kw)
        dt = time.time() - t0
    finally:
        torch.randn = real
        del model.forward
    torch.save(dict(kwargs=kw, init="torch", steps=_records(calls, 1, 3.0, 5.0), out_block=out[:, 480:].clone(),
                    input_checksum=float(cont.double().sum() + noise.double().sum() + spk.double().sum()), ref_seconds=dt),
               os.path.join(GOLD, "dit_base_c5.pt"))
    print(f"dit_base_c5.pt: upstream blockwise sampler {dt:.0f} s", flush=True)


def main():
    which = sys.argv[1:] or ["c4", "c5", "c2"]
    ns = ref_import.load()
    assert ns is not None, "reference not found"
    torch.set_num_threads(os.cpu_count())
    sd = S.synth_state_dict(S.dit_schema(S.ECHO_BASE, blockwise=True), init="torch")
    model = ref_import.build_dit(ns, S.ECHO_BASE, sd)
    with torch.inference_mode():
        for w in which:
            {"c2": gen_c2, "c4": gen_c4, "c5": gen_c5}[w](ns, model)


if __name__ == "__main__":
    main()
