#!/usr/bin/env python
"""One-GPU measurements of the BASELINE.json configs that bench.py's headline line does not cover:
  configs[2] per-GPU share : B independent prompts of the config-2 workload in one batch (--case c3 --batch 8)
  configs[3] long reference: 5-minute speaker reference (6400 latents = 1600 cached speaker patches), Force Speaker
                             speaker_kv_scale=1.5 over speaker_kv_max_layers=24 until t < 0.9, batch 8 (--case c4)
Device-timed generation (sampler graph replay + S1-DAC decode) and, for c4, the once-per-prompt conditioning separately.
Prints one JSON line per case."""
import argparse, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench


def run(case, B, reps):
    from echo_tts_b200 import schema as S
    from echo_tts_b200.inference import load_model_from_state_dict, load_fish_ae_from_state_dict, ae_decode, PCAState
    from echo_tts_b200.sampler import sample_euler_cfg_independent_guidances as sample
    from oracle import echo_oracle as O
    model = load_model_from_state_dict(S.synth_state_dict(S.dit_schema(S.ECHO_BASE, blockwise=False), init="torch"))
    ae = load_fish_ae_from_state_dict(S.synth_ae_state_dict(S.AE_BASE))
    comp, mean, scale = O.synth_pca()
    pca = PCAState(comp.cuda(), mean.cuda(), scale)
    kw = dict(bench.KW)
    t_spk = 212
    if case == "c4":
        t_spk = 6400
        kw.update(speaker_kv_scale=1.5, speaker_kv_max_layers=24, speaker_kv_min_t=0.9)
    bench.T_SPK = t_spk
    ids, mask, spk, smask = [t.cuda() for t in bench.synth_inputs(B)]
    l_spk = t_spk // 4
    S_ = bench.S_LAT
    att = lambda keys: sum(24 * 4 * S_ * k * 2048 for k in keys)
    f_cfg = 3 * S_ * bench.F_TOK + att([S_ + 256 + l_spk, S_ + l_spk, S_ + 256])
    f_plain = S_ * bench.F_TOK + att([S_ + 256 + l_spk])
    flop = B * (20 * f_cfg + 20 * f_plain + bench.FLOP_DECODE)

    def gen(seed):
        lat = sample(model, spk, smask, ids, mask, seed, **kw)
        return ae_decode(ae, pca, lat)

    for i in range(3):
        gen(i)
    clocks = bench.ClockSampler(0)
    clocks.start()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(reps):
        gen(10 + i)
    e1.record()
    torch.cuda.synchronize()
    clocks.stop = True
    ms = e0.elapsed_time(e1) / reps
    # once-per-prompt conditioning alone (text + speaker encoders + 24-layer KV projections)
    ts = []
    for _ in range(3):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record(); model.get_kv_cache_text(ids, mask); model.get_kv_cache_speaker(spk); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    pk = bench.peaks()
    audio = B * bench.AUDIO_SEC_PER_PROMPT
    print(json.dumps(dict(case=case, prompts=B, speaker_latents=t_spk, kwargs={k: v for k, v in kw.items() if v is not None},
                          ms_per_batch=ms, audio_sec_per_gpu_sec=audio / (ms / 1e3), conditioning_ms=sorted(ts)[1],
                          tflops=flop / (ms / 1e3) / 1e12, frac_of_sustained_bf16=flop / (ms / 1e3) / 1e12 / pk["bf16_sustained"],
                          flop_note="sampling + decode, algorithmic (SURVEY 8d); conditioning is inside ms_per_batch but not credited",
                          clocks=clocks.summary())), flush=True)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--case", default="c4", choices=["c3", "c4"])
    ap.add_argument("--batch", type=int, default=8)
    ap.add_argument("--reps", type=int, default=3)
    a = ap.parse_args()
    run(a.case, a.batch, a.reps)
