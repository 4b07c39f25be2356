"""Full-architecture blockwise fixture (BASELINE configs[4] in miniature) produced by EXECUTING the upstream reference:
echo-tts-base architecture incl. the latent encoder, nn.Linear-default-init weights (seed 1234), fp32 on CPU,
16-latent continuation + blocks [16, 32], 4 Euler steps, cfg 3.0 / 5.0 (inference_blockwise.py:161-166), Force-Speaker
scale 1.5 on 12 layers until t < 0.6 (re-applied at every block, inference_blockwise.py:68-70).

    python oracle/gen_golden_blockwise.py        # build container only (needs /root/reference, ~12 GB RAM)
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


def main():
    ns = ref_import.load()
    cfg = S.ECHO_BASE
    sd = S.synth_state_dict(S.dit_schema(cfg), init="torch")
    model = ref_import.build_dit(ns, cfg, sd)
    gen = torch.Generator().manual_seed(21)
    ids, mask = ns.inference.get_text_input_ids_and_mask(["Blockwise generation at full size."], max_length=None)
    spk = torch.randn(1, 212, 80, generator=gen)
    smask = torch.ones(1, 212, dtype=torch.bool)
    cont = torch.randn(1, 16, 80, generator=gen)
    blocks = [16, 32]
    noises = [torch.randn(1, b, 80, generator=gen) for b in blocks]
    it = iter(noises)
    real_randn = torch.randn
    torch.randn = lambda *a, **k: next(it).clone()
    kw = dict(block_sizes=blocks, num_steps=4, cfg_scale_text=3.0, cfg_scale_speaker=5.0, cfg_min_t=0.5, cfg_max_t=1.0,
              truncation_factor=None, rescale_k=None, rescale_sigma=None, speaker_kv_scale=1.5, speaker_kv_max_layers=12,
              speaker_kv_min_t=0.6)
    try:
        t0 = time.time()
        out = ns.inference_blockwise.sample_blockwise_euler_cfg_independent_guidances(
            model, spk, smask, ids, mask, 21, continuation_latent=cont, **kw)
        dt = time.time() - t0
    finally:
        torch.randn = real_randn
    torch.save(dict(ids=ids, text_mask=mask, speaker_latent=spk, speaker_mask=smask, continuation=cont, noises=noises,
                    kwargs=kw, out=out, init="torch", ref_seconds=dt), os.path.join(GOLD, "dit_base_blockwise.pt"))
    print("dit_base_blockwise.pt: reference blockwise sampler seconds:", dt, "out", tuple(out.shape))


if __name__ == "__main__":
    main()
