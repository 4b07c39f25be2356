#!/usr/bin/env python
"""bench.py -- audio-seconds generated per GPU-second on BASELINE.json config 2
(echo-tts-base architecture, random-init, B prompts/GPU x 640 latents x 40 Euler steps, CFG text 3.0 /
speaker 8.0 on t in [0.5, 1], 10 s speaker reference (212 latents), + S1-DAC decode to 44.1 kHz).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--batch B] [--impl reference]

One "step" = one complete generation (conditioning KV build + 40 sampler steps + decode + crop index) of B
prompts per GPU.  Prints ONE JSON line on rank 0 (contract in the task prompt).  `--impl reference` times the
reference algorithm's CPU path (the oracle restatement: the reference itself cannot travel to the GPU box).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

S_LAT, N_STEPS, T_SPK, L_TEXT, N_TEXT_VALID = 640, 40, 212, 768, 256
AUDIO_SEC_PER_PROMPT = S_LAT * 2048 / 44100.0
KW = dict(num_steps=N_STEPS, cfg_scale_text=3.0, cfg_scale_speaker=8.0, cfg_min_t=0.5, cfg_max_t=1.0,
          truncation_factor=None, rescale_k=None, rescale_sigma=None, speaker_kv_scale=None,
          speaker_kv_max_layers=None, speaker_kv_min_t=None, sequence_length=S_LAT)
# algorithmic FLOPs (SURVEY.md 8d / BASELINE.md 4)
F_TOK = 2 * 24 * (5 * 2048 ** 2 + 3 * 2048 * 5888) + 4 * 80 * 2048


def attn_flops(rows_valid_keys):
    return sum(24 * 4 * S_LAT * lv * 2048 for lv in rows_valid_keys)


L_SPK = T_SPK // 4
FLOP_CFG = 3 * S_LAT * F_TOK + attn_flops([S_LAT + N_TEXT_VALID + L_SPK, S_LAT + L_SPK, S_LAT + N_TEXT_VALID])
FLOP_PLAIN = S_LAT * F_TOK + attn_flops([S_LAT + N_TEXT_VALID + L_SPK])
FLOP_DECODE = 4.342e12
FLOP_COND = 0.26e12
FLOP_GEN = 20 * FLOP_CFG + 20 * FLOP_PLAIN + FLOP_DECODE + FLOP_COND


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(p):
        d = json.load(open(p))
        return dict(bf16=d["bf16_tflops"], bf16_sustained=d["bf16_tflops_sustained"], hbm=d["hbm_gbs"], src="measured")
    return dict(bf16=1590.0, bf16_sustained=1400.0, hbm=6650.0, src="fallback")


def synth_inputs(B, seed=7):
    g = torch.Generator().manual_seed(seed)
    ids = torch.zeros((B, L_TEXT), dtype=torch.int32)
    mask = torch.zeros((B, L_TEXT), dtype=torch.bool)
    ids[:, 1:N_TEXT_VALID] = torch.randint(32, 127, (B, N_TEXT_VALID - 1), generator=g, dtype=torch.int32)
    mask[:, :N_TEXT_VALID] = True
    spk = torch.randn((B, T_SPK, 80), generator=g)
    smask = torch.ones((B, T_SPK), dtype=torch.bool)
    return ids, mask, spk, smask


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop = index, [], False

    def run(self):
        while not self.stop:
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i",
                                      str(self.index)], capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([c.strip() for c in out.split(",")])
            except Exception:
                pass
            time.sleep(0.1)

    def summary(self):
        sm = sorted(float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i] == "Active" for r in self.rows)]
        mx = max((float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()), default=None)
        return dict(sm_mhz=(sm[len(sm) // 2] if sm else None), sm_max_mhz=mx, reasons=reasons, samples=len(sm))


def cpu_reference_sample(threads=None, repeats=1):
    """Time the reference algorithm's CPU path (oracle port, fp32, all host threads) on a bounded sample of the
    workload: ONE CFG forward of config 2 (3 rows x 640 latents, 768-padded text with 256 valid, 53 speaker
    patches), then scale by algorithmic FLOPs to the whole generation."""
    from echo_tts_b200 import schema as S
    from oracle import echo_oracle as O
    threads = threads or os.cpu_count()
    torch.set_num_threads(threads)
    cfg = S.ECHO_BASE
    sd = S.synth_state_dict(S.dit_schema(cfg, blockwise=False), init="torch")
    ids, mask, spk, smask = synth_inputs(1)
    with torch.inference_mode():
        kv_t = O.kv_cache_text(sd, cfg, ids, mask)
        kv_s = O.kv_cache_speaker(sd, cfg, spk)
        x = torch.randn(1, S_LAT, 80, generator=torch.Generator().manual_seed(1))
        tm3 = torch.cat([mask, torch.zeros_like(mask), mask], 0)
        sm3 = torch.cat([smask, smask, torch.zeros_like(smask)], 0)
        rep3 = lambda kv: [(torch.cat([k] * 3, 0), torch.cat([v] * 3, 0)) for k, v in kv]
        kv_t3, kv_s3 = rep3(kv_t), rep3(kv_s)
        times = []
        for _ in range(repeats):
            t0 = time.perf_counter()
            O.dit_forward(sd, cfg, torch.cat([x] * 3, 0), torch.ones(3) * 0.9, tm3, sm3, kv_t3, kv_s3)
            times.append(time.perf_counter() - t0)
    sec = sorted(times)[len(times) // 2]
    gen_sec = sec * FLOP_GEN / FLOP_CFG
    return dict(value=AUDIO_SEC_PER_PROMPT / gen_sec, unit="audio-sec/CPU-sec", cores=threads, kind="port",
                sample=f"1 CFG forward (3x640 tokens, {FLOP_CFG / 1e12:.2f} TFLOP) in {sec:.2f} s, median of {repeats}; "
                       f"scaled by FLOPs to the {FLOP_GEN / 1e12:.1f}-TFLOP generation"), sec


def eager_gpu_port(reps=3):
    """The practical bar of SURVEY 8d / BASELINE.md: the reference ALGORITHM as plain PyTorch eager ops on the same B200
    (the oracle's functional restatement executed under torch.device("cuda"); the upstream repo itself cannot travel to
    the GPU box), at the reference's default dtypes -- bf16 DiT (inference.py:17), fp32 S1-DAC (inference.py:83).  One
    three-branch CFG forward, one plain forward and one decode are timed (median of `reps` after a warm-up) and summed
    20 / 20 / 1 as in config 2; conditioning and the reference's per-step host syncs are left out (in its favour)."""
    from echo_tts_b200 import schema as S
    from oracle import echo_oracle as O
    cfg = S.ECHO_BASE
    bf = torch.bfloat16
    sd = {k: (v.cuda().to(bf) if v.is_floating_point() else v.cuda())
          for k, v in S.synth_state_dict(S.dit_schema(cfg, blockwise=False), init="torch").items()}
    asd = {k: v.cuda() for k, v in S.synth_ae_state_dict(S.AE_BASE).items()}
    comp, mean, scale = [t.cuda() if torch.is_tensor(t) else t for t in O.synth_pca()]
    ids, mask, spk, smask = [t.cuda() for t in synth_inputs(1)]

    def timed(fn):
        fn()
        ts = []
        for _ in range(reps):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            a.record(); fn(); b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        return sorted(ts)[reps // 2]

    with torch.inference_mode(), torch.device("cuda"):
        kv_t = O.kv_cache_text(sd, cfg, ids, mask)
        kv_s = O.kv_cache_speaker(sd, cfg, spk.to(bf))
        x = torch.randn(1, S_LAT, 80).to(bf)
        tm3 = torch.cat([mask, torch.zeros_like(mask), mask], 0)
        sm3 = torch.cat([smask, smask, torch.zeros_like(smask)], 0)
        rep3 = lambda kv: [(torch.cat([k] * 3, 0), torch.cat([v] * 3, 0)) for k, v in kv]   # inference.py:489-492
        kv_t3, kv_s3 = rep3(kv_t), rep3(kv_s)
        cfg_ms = timed(lambda: O.dit_forward(sd, cfg, torch.cat([x] * 3, 0), torch.ones(3).to(bf) * 0.9, tm3, sm3, kv_t3, kv_s3))
        plain_ms = timed(lambda: O.dit_forward(sd, cfg, x, torch.ones(1).to(bf) * 0.4, mask, smask, kv_t, kv_s))
        z = torch.randn(1, S_LAT, 80)
        dec_ms = timed(lambda: O.ae_decode(asd, S.AE_BASE, comp, mean, scale, z))
    gen_ms = 20 * cfg_ms + 20 * plain_ms + dec_ms
    return dict(value=AUDIO_SEC_PER_PROMPT / (gen_ms / 1e3), unit="audio-sec/sec", kind="port",
                what="oracle restatement as torch eager ops on cuda:0, bf16 DiT + fp32 S1-DAC (the reference's default "
                     "dtypes); 20 x CFG forward + 20 x plain forward + decode, device-timed",
                cfg_forward_ms=cfg_ms, plain_forward_ms=plain_ms, decode_ms=dec_ms, generation_ms=gen_ms)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cb, sec = cpu_reference_sample(repeats=max(1, min(args.steps, 3)))
    line = dict(metric="audio-sec generated per GPU-sec, 640-latent 40-step CFG", impl="reference", value=cb["value"],
                unit="audio-sec/sec", n_gpus=args.gpus, steps=args.steps, warmup=args.warmup,
                ms_per_step=sec * FLOP_GEN / FLOP_CFG * 1e3, higher_is_better=True, scaling="weak", vs_baseline=None,
                dtype="f32", data="synthetic", config=workload_config(1), cpu_baseline=cb,
                e2e=dict(value=cb["value"], unit="audio-sec/sec", h2d_bytes_per_step=0, d2h_bytes_per_step=0))
    print(json.dumps(line), flush=True)


def workload_config(B):
    return dict(workload="BASELINE.json configs[1]: echo-tts-base arch random-init, 640 latents x 40 Euler steps, "
                         "CFG text 3.0/speaker 8.0 on t in [0.5,1] (20 three-branch + 20 single forwards), 212-latent "
                         "speaker ref, text 768-padded (256 valid), + S1-DAC decode to 1,310,720 samples",
                prompts_per_gpu=B, l2="weights (5.6 GB bf16) exceed the 126 MB L2; no explicit flush")


def gemm_roofline(model, pk):
    """Dominant kernel: the CTA-pair tcgen05 GEMM (csrc/gemm2.cu).  Timed live with CUDA events on the launch stream over all 24 layers'
    w1|w3 operands (24 x 48 MB > L2) at the CFG-step shape M=1920, N=11776, K=2048 (43 % of step FLOPs)."""
    from echo_tts_b200 import ops
    M, D, F2 = 3 * S_LAT, 2048, 2 * 5888
    a = torch.randn((M, D), device="cuda").to(torch.bfloat16)
    out = torch.empty((M, F2 // 2), device="cuda", dtype=torch.bfloat16)
    for lw in model.layers[:3]:
        ops.gemm_swiglu(a, lw["w13"], out)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 3
    e0.record()
    for _ in range(reps):
        for lw in model.layers:
            ops.gemm_swiglu(a, lw["w13"], out)
    e1.record()
    torch.cuda.synchronize()
    n = reps * len(model.layers)
    sec = e0.elapsed_time(e1) / 1e3 / n
    flops = 2.0 * M * F2 * D
    ach = flops / sec / 1e12
    # traffic: dram__bytes_read.sum + dram__bytes_write.sum of this launch in the committed `ncu --set full` capture
    # (profiles/r1_ncu_hot_kernels.md: 56.17 MB + 5.18 MB), per launch like `achieved`
    return dict(bound="tensor", kernel="gemm_pair_kernel<SWIGLU> (tcgen05 cta_group::2, 256x256 tiles) M=1920 N=11776 K=2048",
                achieved=ach, peak=pk["bf16"], unit="TFLOP/s", frac=ach / pk["bf16"], traffic=61.35e6,
                traffic_unit="bytes/launch (ncu dram read+write)", flop_per_launch=flops,
                peak_source=pk["src"] + " burst", launches_timed=n, us_per_launch=sec * 1e6,
                note="main loop is shared-memory-bandwidth bound: 669 clk per 64-wide k-block = 80 KB of operand reads + TMA "
                     "writes per SM at 120 of 128 B/clk; MMAs alone 491 clk (DESIGN.md section 4, tools/gemm_trace.py)")


def config5_streaming(model, ae, pca, dev, reps=3):
    """BASELINE.json configs[4]: 4 blocks of 160 latents with prefix conditioning (inference_blockwise.py:14-123, its
    __main__ settings :161-166: cfg 3.0 / 5.0), each block decoded by the S1-DAC decode stream as soon as it is sampled
    and its audio copied to the host.  Wall clock per request, median of `reps` after one warm-up request."""
    from echo_tts_b200.inference_blockwise import stream_blockwise_audio
    kw = dict({k: v for k, v in KW.items() if k != "sequence_length"}, cfg_scale_speaker=5.0)
    blocks = [160] * 4

    def request(incremental):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        marks, n = [], 0
        for wav, a, b in stream_blockwise_audio(model, ae, pca, dev[2], dev[3], dev[0], dev[1], 0, blocks,
                                                incremental_prefix=incremental, **kw):
            n += wav.cpu().numel()                        # D2H of the chunk = the moment a client could play it
            marks.append((time.perf_counter() - t0) * 1e3)
        return marks, n

    rec = {}
    for name, inc in (("incremental_prefix", True), ("recomputed_prefix", False)):
        request(inc)
        runs = sorted((request(inc) for _ in range(reps)), key=lambda r: r[0][-1])
        marks, n = runs[reps // 2]
        rec[name] = dict(total_ms=marks[-1], first_audio_ms=marks[0], block_done_ms=marks, samples_out=n,
                         audio_sec_per_sec=n / 44100 / (marks[-1] / 1e3))
    rec["what"] = ("stream_blockwise_audio(): 4 x 160 latents x 40 steps, cfg 3.0/5.0 on t in [0.5,1], per-block streaming "
                   "S1-DAC decode + D2H, " + str(dev[0].shape[0]) + " prompt(s), eager launches; recomputed_prefix = the reference's full prefix re-encode per block")
    return rec


def run_gpu(args):
    from echo_tts_b200 import schema as S, _lib
    from echo_tts_b200.inference import (load_model_from_state_dict, load_fish_ae_from_state_dict, ae_decode, PCAState,
                                         find_flattening_point)
    from echo_tts_b200.sampler import sample_euler_cfg_independent_guidances as sample
    from oracle import echo_oracle as O   # only synth_pca() (synthetic PCA constants) and the cpu_baseline leg
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    B = args.batch
    pk = peaks()
    t0 = time.time()
    model = load_model_from_state_dict(S.synth_state_dict(S.dit_schema(S.ECHO_BASE, blockwise=True), init="torch"))
    ae = load_fish_ae_from_state_dict(S.synth_ae_state_dict(S.AE_BASE, encode=True))
    comp, mean, scale = O.synth_pca()
    pca = PCAState(comp.cuda(), mean.cuda(), scale)
    load_s = time.time() - t0
    ids, mask, spk, smask = synth_inputs(B, seed=7 + rank)
    host = [t.pin_memory() for t in (ids, mask, spk, smask)]
    dev = [t.cuda() for t in host]
    lib = _lib.load()

    def generate(inp, seed):
        lat = sample(model, inp[2], inp[3], inp[0], inp[1], seed, **KW)
        wav = ae_decode(ae, pca, lat)
        return wav, lat

    def generate_e2e(seed):
        inp = [t.cuda(non_blocking=True) for t in host]               # H2D of this step's inputs (pinned)
        wav, lat = generate(inp, seed)
        fp = [find_flattening_point(lat[b]) for b in range(B)]         # D2H of the crop indices
        return wav.cpu(), fp                                           # D2H of the waveform

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            import torch.distributed as dist
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, n):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(n):
            fn(i)
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
        if world > 1:
            import torch.distributed as dist
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    lib.echo_reset_launch_count()
    generate(dev, 0)                                                   # eager: counts launches of one generation
    launches_per_gen = int(lib.echo_launch_count())
    for i in range(max(args.warmup, 3)):
        generate(dev, 1 + i)
        generate_e2e(1 + i)
    clocks = ClockSampler(local)
    clocks.start()
    ms = timed(lambda i: generate(dev, 100 + i), args.steps)
    ms_e2e = timed(lambda i: generate_e2e(200 + i), args.steps)
    clocks.stop = True
    # phase breakdown of one generation (events on the launch stream)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    torch.cuda.synchronize()
    ev[0].record(); lat = sample(model, dev[2], dev[3], dev[0], dev[1], 5, **KW); ev[1].record()
    ae_decode(ae, pca, lat); ev[2].record(); torch.cuda.synchronize()
    sampler_ms, decode_ms = ev[0].elapsed_time(ev[1]), ev[1].elapsed_time(ev[2])
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return
    audio = world * B * AUDIO_SEC_PER_PROMPT * args.steps
    value, e2e = audio / (ms / 1e3), audio / (ms_e2e / 1e3)

    def sampler_ms_median(kw, n=5):
        """median device time of one graph-replayed sample() call under schedule `kw`"""
        for _ in range(3):                                             # eager, capture, replay
            sample(model, dev[2], dev[3], dev[0], dev[1], 0, **kw)
        ts = []
        for _ in range(n):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            a.record(); sample(model, dev[2], dev[3], dev[0], dev[1], 0, **kw); b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        return sorted(ts)[n // 2]

    # the whole reference-facing pipeline from a 10 s waveform: S1-DAC encode of the speaker clip, text
    # tokenisation, conditioning, 40 Euler steps, decode, flattening-point crop; host waveform in, host waveform out
    from functools import partial
    from echo_tts_b200.inference import sample_pipeline
    wave = ((torch.rand(1, 441000, generator=torch.Generator().manual_seed(11)) * 2 - 1) * 0.5).pin_memory()
    sample_fn = partial(sample, **KW)
    text = "[S1] " + ("the quick brown fox jumps over the lazy dog " * 6)[:250]        # 255 bytes + BOS = 256 valid ids

    def pipeline():
        audio, _ = sample_pipeline(model, ae, pca, sample_fn, text, wave, 0, pad_to_max_text_length=768)
        return audio.cpu()
    pipe_ms = []
    for i in range(6):                                                 # first calls warm the encoder and capture the graph
        t_p = time.perf_counter()
        out_wave = pipeline()
        torch.cuda.synchronize()
        pipe_ms.append((time.perf_counter() - t_p) * 1e3)
    pipeline_rec = dict(ms=sorted(pipe_ms[3:])[1], samples_out=int(out_wave.shape[-1]), speaker_clip_samples=441000,
                        what="sample_pipeline(): host waveform -> S1-DAC encode -> sampler -> decode -> crop -> host, "
                             "wall clock, median of 3 after 3 warm-up calls")

    c5 = config5_streaming(model, ae, pca, dev) if (world == 1 and not args.no_config5) else None

    # per-step latency: slope between 40- and 20-step schedules, all steps guided / none guided
    all_cfg, no_cfg = dict(KW, cfg_min_t=0.0, cfg_max_t=1.0), dict(KW, cfg_min_t=2.0, cfg_max_t=3.0)
    step_ms = {name: (sampler_ms_median(dict(kw, num_steps=40)) - sampler_ms_median(dict(kw, num_steps=20))) / 20
               for name, kw in (("cfg_step", all_cfg), ("plain_step", no_cfg))}
    step_ms["method"] = "(40-step - 20-step sample() graph replay, median of 5) / 20, all steps guided / none guided"
    roof = gemm_roofline(model, pk)
    step_tflops = B * FLOP_GEN / (ms / 1e3 / args.steps) / 1e12
    line = dict(metric="audio-sec generated per GPU-sec, 640-latent 40-step CFG", value=value, unit="audio-sec/sec",
                n_gpus=world, steps=args.steps, warmup=max(args.warmup, 3), ms_per_step=ms / args.steps,
                higher_is_better=True, scaling="weak", vs_baseline=None, dtype="bf16", data="synthetic",
                config=workload_config(B), clocks=clocks.summary(),
                e2e=dict(value=e2e, unit="audio-sec/sec",
                         h2d_bytes_per_step=sum(t.numel() * t.element_size() for t in host),
                         d2h_bytes_per_step=B * S_LAT * 2048 * 4 + 4 * B, ms_per_step=ms_e2e / args.steps),
                gpu_launches=launches_per_gen * args.steps, launches_per_generation=launches_per_gen,
                roofline=roof,
                generation_roofline=dict(bound="tensor", achieved=step_tflops, peak=pk["bf16_sustained"], unit="TFLOP/s",
                                         frac=step_tflops / pk["bf16_sustained"], flop_per_generation=B * FLOP_GEN,
                                         peak_source=pk["src"] + " sustained"),
                breakdown_ms=dict(conditioning_plus_sampler=sampler_ms, decode=decode_ms), step_latency_ms=step_ms, pipeline_from_waveform=pipeline_rec,
                load_seconds=load_s)
    if c5 is not None:
        line["config5_streaming"] = c5
    if world == 1 and B == 1 and not args.no_eager_gpu:
        del model, ae
        torch.cuda.empty_cache()
        try:
            line["eager_torch_gpu"] = eager_gpu_port()
            line["eager_torch_gpu"]["speedup_of_this_engine"] = value / line["eager_torch_gpu"]["value"]
        except Exception as e:                                         # a baseline must never cost the bench line
            line["eager_torch_gpu"] = dict(error=f"{type(e).__name__}: {e}"[:300])
        model = ae = None
    if not args.no_cpu_baseline:
        del model, ae
        torch.cuda.empty_cache()
        line["cpu_baseline"], _ = cpu_reference_sample()
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=1, help="prompts per GPU")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-eager-gpu", action="store_true", help="skip the torch-eager-on-B200 baseline record")
    ap.add_argument("--no-config5", action="store_true", help="skip the BASELINE configs[4] streaming record")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    run_gpu(args)


if __name__ == "__main__":
    main()
