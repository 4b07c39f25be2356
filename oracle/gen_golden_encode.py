#!/usr/bin/env python
"""Golden vectors for the S1-DAC *encode* path, produced by EXECUTING the upstream reference
(imported read-only from /root/reference; build container only) on the deterministic synthetic weights of
``echo_tts_b200/schema.py``.  Test infrastructure, never imported by the product package.

    python oracle/gen_golden_encode.py

  tests/golden/ae_encode_tiny.pt   AE_TINY dims (reference classes instantiated at those dims), 2 ragged clips long
                                   enough to exercise both attention windows (640 encoder frames > 512, 160 > 16)
  tests/golden/ae_encode_base.pt   the real build_ae() dims, one short ragged clip (12 frames) + the PCA projection
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from echo_tts_b200 import schema as S  # noqa: E402
from oracle import echo_oracle as O  # noqa: E402
from oracle import ref_import  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def main():
    ns = ref_import.load()
    assert ns is not None, "reference not found"
    half = lambda d: {k: (v.half() if v.is_floating_point() else v) for k, v in d.items()}   # fp16 storage, small files
    g = torch.Generator().manual_seed(21)
    cfg = S.AE_TINY
    ae = ref_import.build_ae_cfg(ns, cfg, S.synth_ae_state_dict(cfg, encode=True))
    audio = (torch.rand(2, 1, cfg.hop * 160 - 37, generator=g) * 2 - 1) * 0.8
    st = ref_import.encode_stages(ae, audio)
    torch.save(dict(audio=audio, codes=st["codes"], **half({k: st[k] for k in ("z_enc", "z_pre", "z_q")})),
               os.path.join(GOLD, "ae_encode_tiny.pt"))

    cfg = S.AE_BASE
    ae = ref_import.build_ae_cfg(ns, cfg, S.synth_ae_state_dict(cfg, encode=True))
    audio = (torch.rand(1, 1, cfg.hop * 12 - 1000, generator=g) * 2 - 1) * 0.8
    st = ref_import.encode_stages(ae, audio)
    comp, mean, scale = O.synth_pca()
    pca = ns.inference.PCAState(pca_components=comp, pca_mean=mean, latent_scale=scale)
    lat = ns.inference.ae_encode(ae, pca, audio)
    torch.save(dict(audio=audio, codes=st["codes"], latent=lat, **half({k: st[k] for k in ("z_enc", "z_pre", "z_q")})),
               os.path.join(GOLD, "ae_encode_base.pt"))
    for n in ("ae_encode_tiny.pt", "ae_encode_base.pt"):
        print(n, os.path.getsize(os.path.join(GOLD, n)) // 1024, "KiB")


if __name__ == "__main__":
    main()
