#!/usr/bin/env python
"""Profiling driver for the streaming path (BASELINE configs[4]): ONE incremental prefix-KV append of a 160-latent block
(echo_encode_kv, past = 40 patches) and ONE 160-frame S1-DAC decode-stream step (echo_ae_decode_stream_step, window
already full), launched between cudaProfilerStart/Stop so that `ncu --profile-from-start off` sees exactly these launches.

    python tools/profile_stream.py
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402


def main():
    from echo_tts_b200 import schema as S
    from echo_tts_b200.inference import load_model_from_state_dict, load_fish_ae_from_state_dict, PCAState
    from echo_tts_b200.streaming import PrefixKVStream
    from oracle import echo_oracle as O   # synthetic PCA constants only
    model = load_model_from_state_dict(S.synth_state_dict(S.dit_schema(S.ECHO_BASE, blockwise=True), init="torch"))
    ae = load_fish_ae_from_state_dict(S.synth_ae_state_dict(S.AE_BASE))
    comp, mean, scale = O.synth_pca()
    pca = PCAState(comp.cuda(), mean.cuda(), scale)
    lat = torch.randn(1, 480, 80, device="cuda", generator=torch.Generator(device="cuda").manual_seed(3))
    ps = PrefixKVStream(model, 1, 640)
    st = ae.decode_stream()
    for i in (0, 160):                        # warm-up: two blocks (the attention window of the decoder fills up)
        ps.append(lat[:, i:i + 160])
        st.step_latent(lat[:, i:i + 160], pca.pca_components, pca.pca_mean, pca.latent_scale)
    torch.cuda.synchronize()
    torch.cuda.profiler.start()
    ps.append(lat[:, 320:480])
    st.step_latent(lat[:, 320:480], pca.pca_components, pca.pca_mean, pca.latent_scale)
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
    print("profile_stream ok")


if __name__ == "__main__":
    main()
