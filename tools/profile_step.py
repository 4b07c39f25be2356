#!/usr/bin/env python
"""Profiling driver: conditioning + ONE three-branch CFG step + ONE plain step + S1-DAC decode at BASELINE config 2
shapes (S=640, text 768-padded / 256 valid, 212-latent speaker ref), launched eagerly (no CUDA graph) between
cudaProfilerStart/Stop so that `ncu --profile-from-start off` sees exactly these launches.

    python tools/profile_step.py [--batch B] [--no-decode]
"""
import argparse
import os
import sys

os.environ["ECHO_CUDA_GRAPH"] = "0"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=1)
    ap.add_argument("--no-decode", action="store_true")
    args = ap.parse_args()
    from echo_tts_b200 import schema as S
    from echo_tts_b200.inference import load_model_from_state_dict, load_fish_ae_from_state_dict, ae_decode, PCAState
    from echo_tts_b200.sampler import sample_euler_cfg_independent_guidances as sample
    model = load_model_from_state_dict(S.synth_state_dict(S.dit_schema(S.ECHO_BASE, blockwise=False), init="torch"))
    ae = load_fish_ae_from_state_dict(S.synth_ae_state_dict(S.AE_BASE))
    comp, mean, scale = S.synth_pca()
    pca = PCAState(comp.cuda(), mean.cuda(), scale)
    ids, mask, spk, smask = [t.cuda() for t in bench.synth_inputs(args.batch)]
    kw = dict(bench.KW, num_steps=2)          # t = 0.999 (CFG on), 0.4995 (CFG off)

    def run():
        lat = sample(model, spk, smask, ids, mask, 0, **kw)
        if not args.no_decode:
            ae_decode(ae, pca, lat)
    run()
    torch.cuda.synchronize()
    torch.cuda.profiler.start()
    run()
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
    print("profile_step ok")


if __name__ == "__main__":
    main()
