#!/usr/bin/env python
"""S1-DAC encode on the GPU: parity numbers against the committed reference fixtures and the time of one full
30-second chunk (640 frames = 1 310 720 samples, the unit get_speaker_latent_and_mask encodes)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from echo_tts_b200 import schema as S  # noqa: E402
from echo_tts_b200.inference import load_fish_ae_from_state_dict  # noqa: E402


def rel(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm())


def main():
    for name, cfg in (("ae_encode_tiny.pt", S.AE_TINY), ("ae_encode_base.pt", S.AE_BASE)):
        g = torch.load(os.path.join(ROOT, "tests", "golden", name))
        dac = load_fish_ae_from_state_dict(S.synth_ae_state_dict(cfg, encode=True), cfg=cfg)
        st = dac.encode_stages(g["audio"].cuda())
        same = (st["codes"].cpu() == g["codes"])
        print(f"{name}: z_enc {rel(st['z_enc'].cpu(), g['z_enc'].float()):.2e}  z_pre {rel(st['z_pre'].cpu(), g['z_pre'].float()):.2e}  "
              f"z_q {rel(st['z_q'].cpu(), g['z_q'].float()):.2e}  level-0 codes equal {same[:, 0].float().mean():.3f}  "
              f"all-level frames equal {same.all(1).float().mean():.3f}")
    cfg = S.AE_BASE
    audio = (torch.rand(1, 1, 640 * cfg.hop, generator=torch.Generator().manual_seed(0)) * 2 - 1).cuda()
    for _ in range(2):
        dac.encode_zq(audio)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        dac.encode_zq(audio)
    e1.record()
    torch.cuda.synchronize()
    print(f"encode_zq of one 30 s chunk (1 310 720 samples -> 640 frames): {e0.elapsed_time(e1) / 3:.2f} ms; "
          f"peak memory {torch.cuda.max_memory_allocated() / 2**30:.2f} GiB")


if __name__ == "__main__":
    main()
