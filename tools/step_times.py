#!/usr/bin/env python
"""Where a generation's time goes (CUDA events, config-2 shapes): conditioning encoders, modulation table, sampler
steps (graph replay) split into CFG / plain steps by running all-CFG and no-CFG schedules, decode."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402


def timeit(fn, n=3):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def main():
    from echo_tts_b200 import schema as S
    from echo_tts_b200.inference import load_model_from_state_dict, load_fish_ae_from_state_dict, ae_decode, PCAState
    from echo_tts_b200.sampler import sample_euler_cfg_independent_guidances as sample, StepPlan
    from oracle import echo_oracle as O
    model = load_model_from_state_dict(S.synth_state_dict(S.dit_schema(S.ECHO_BASE, blockwise=False), init="torch"))
    ae = load_fish_ae_from_state_dict(S.synth_ae_state_dict(S.AE_BASE))
    comp, mean, scale = O.synth_pca()
    pca = PCAState(comp.cuda(), mean.cuda(), scale)
    ids, mask, spk, smask = [t.cuda() for t in bench.synth_inputs(1)]
    print(f"text KV cache     {timeit(lambda: model.get_kv_cache_text(ids, mask)):8.3f} ms")
    print(f"speaker KV cache  {timeit(lambda: model.get_kv_cache_speaker(spk)):8.3f} ms")
    plan = StepPlan(model.device, 40, 0.5, 1.0, None, None, None, 24)
    print(f"modulation table  {timeit(lambda: model.build_mod_table(plan.t_dev)):8.3f} ms")
    for name, kw in (("config 2 (20 CFG + 20 plain)", dict(bench.KW)), ("40 CFG steps", dict(bench.KW, cfg_min_t=0.0)),
                     ("40 plain steps", dict(bench.KW, cfg_min_t=2.0, cfg_max_t=3.0))):
        for _ in range(3):
            sample(model, spk, smask, ids, mask, 0, **kw)       # eager, capture, replay
        ms = timeit(lambda: sample(model, spk, smask, ids, mask, 0, **kw))
        print(f"sample() {name:30s} {ms:8.3f} ms")
    lat = sample(model, spk, smask, ids, mask, 0, **bench.KW)
    print(f"ae_decode         {timeit(lambda: ae_decode(ae, pca, lat)):8.3f} ms")


if __name__ == "__main__":
    main()
