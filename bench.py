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
    """Synthetic request of the configs[1] shapes (shared with the parity tests and the fixture generators)."""
    from echo_tts_b200 import schema as S
    return S.synth_request(B, n_valid=N_TEXT_VALID, pad_to=L_TEXT, t_spk=T_SPK, seed=seed)


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
    workload, as BASELINE.md section 5 lays out: the conditioning KV build, ONE three-branch CFG forward, ONE single
    forward and the whole S1-DAC decode of 640 latents are each timed at the configs[1] shapes; a generation is
    conditioning + 20 CFG forwards + 20 single forwards + decode (scaled by STEP COUNT only, never by FLOPs)."""
    from echo_tts_b200 import schema as S
    from oracle import echo_oracle as O
    threads = threads or os.cpu_count()
    torch.set_num_threads(threads)
    cfg = S.ECHO_BASE
    sd = S.synth_state_dict(S.dit_schema(cfg, blockwise=False), init="torch")
    asd = S.synth_ae_state_dict(S.AE_BASE)
    comp, mean, scale = S.synth_pca()
    ids, mask, spk, smask = synth_inputs(1)
    med = lambda ts: sorted(ts)[len(ts) // 2]

    def clock(fn, n):
        ts, out = [], None
        for _ in range(n):
            t0 = time.perf_counter()
            out = fn()
            ts.append(time.perf_counter() - t0)
        return med(ts), out
    with torch.inference_mode():
        cond_s, (kv_t, kv_s) = clock(lambda: (O.kv_cache_text(sd, cfg, ids, mask), O.kv_cache_speaker(sd, cfg, spk)), 1)
        x = torch.randn(1, S_LAT, 80, generator=torch.Generator().manual_seed(1))
        tm3 = torch.cat([mask, torch.zeros_like(mask), mask], 0)
        sm3 = torch.cat([smask, smask, torch.zeros_like(smask)], 0)
        rep3 = lambda kv: [(torch.cat([k] * 3, 0), torch.cat([v] * 3, 0)) for k, v in kv]
        kv_t3, kv_s3 = rep3(kv_t), rep3(kv_s)
        cfg_s, _ = clock(lambda: O.dit_forward(sd, cfg, torch.cat([x] * 3, 0), torch.ones(3) * 0.9, tm3, sm3, kv_t3, kv_s3), repeats)
        plain_s, _ = clock(lambda: O.dit_forward(sd, cfg, x, torch.ones(1) * 0.4, mask, smask, kv_t, kv_s), repeats)
        del sd, kv_t3, kv_s3
        dec_s, _ = clock(lambda: O.ae_decode(asd, S.AE_BASE, comp, mean, scale, x), 1)
    gen_sec = cond_s + 20 * cfg_s + 20 * plain_s + dec_s
    return dict(value=AUDIO_SEC_PER_PROMPT / gen_sec, unit="audio-sec/CPU-sec", cores=threads, kind="port",
                sample=f"configs[1] shapes, fp32, {threads} threads: conditioning {cond_s:.2f} s, one 3-branch CFG forward "
                       f"{cfg_s:.2f} s, one single forward {plain_s:.2f} s (median of {repeats}), S1-DAC decode of 640 latents "
                       f"{dec_s:.2f} s; generation = conditioning + 20 x CFG + 20 x single + decode = {gen_sec:.1f} s "
                       f"(step-count scaling only)",
                conditioning_s=cond_s, cfg_forward_s=cfg_s, plain_forward_s=plain_s, decode_s=dec_s,
                generation_s=gen_sec), gen_sec


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
    comp, mean, scale = [t.cuda() if torch.is_tensor(t) else t for t in S.synth_pca()]
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
                ms_per_step=sec * 1e3, higher_is_better=True, scaling="weak", vs_baseline=None,
                dtype="f32", data="synthetic", config=workload_config(1), cpu_baseline=cb,
                e2e=dict(value=cb["value"], unit="audio-sec/sec", h2d_bytes_per_step=0, d2h_bytes_per_step=0))
    print(json.dumps(line), flush=True)


def workload_config(B):
    return dict(workload="BASELINE.json configs[1]: echo-tts-base arch random-init, 640 latents x 40 Euler steps, "
                         "CFG text 3.0/speaker 8.0 on t in [0.5,1] (20 three-branch + 20 single forwards), 212-latent "
                         "speaker ref, text 768-padded (256 valid), + S1-DAC decode to 1,310,720 samples",
                prompts_per_gpu=B, l2="weights (5.6 GB bf16) exceed the 126 MB L2; no explicit flush")


def _event_us(fn, n, warm=2):
    """Mean microseconds of fn(i) over n back-to-back launches replayed from a CUDA graph (no Python / ctypes launch
    overhead between kernels), CUDA events on the launch stream, after `warm` eager calls and one untimed replay."""
    for i in range(warm):
        fn(i)
    torch.cuda.synchronize()
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr):
        for i in range(n):
            fn(i)
    gr.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    gr.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / n


def roofline_traffic():
    """`roofline.traffic`: dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel, taken from the
    committed `ncu --set full` capture named in profiles/roofline_traffic.json (written by tools/ncu_summary.py); null
    when no capture of the current kernel is committed -- never a literal in this file."""
    p = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if not os.path.isfile(p):
        return None, None
    d = json.load(open(p))
    return d.get("bytes_per_launch"), d.get("source")


def gemm_roofline(model, pk):
    """Dominant kernel: the CTA-pair tcgen05 GEMM (csrc/gemm2.cu).  Timed live with CUDA events on the launch stream over all 24 layers'
    w1|w3 operands (24 x 48 MB > L2) at the CFG-step shape M=1920, N=11776, K=2048 (43 % of step FLOPs).
    `cublas_same_shape`: every projection of a decoder layer at both step shapes (M = 1920 guided, 640 un-guided) through
    this library's kernel WITH its fused epilogue and through torch.matmul (cuBLAS, plain bf16 GEMM, no epilogue), same
    event timing, each cycling the 24 layers' weights so operands stream from HBM."""
    from echo_tts_b200 import ops, _lib as L
    D, F, H = 2048, 5888, 16
    layers = model.layers
    nl = len(layers)
    table = {}
    for M in (3 * S_LAT, S_LAT):
        R = M // S_LAT
        xn = torch.randn((M, D), device="cuda").to(torch.bfloat16)
        hm = torch.randn((M, F), device="cuda").to(torch.bfloat16)
        x = torch.randn((M, D), device="cuda")
        gate = torch.randn((R, D), device="cuda")
        qkvg = torch.empty((4, M, D), device="cuda", dtype=torch.bfloat16)
        hmid = torch.empty((M, F), device="cuda", dtype=torch.bfloat16)
        ns = model.k_splits(M)
        parts = torch.empty((ns, M, D), device="cuda") if ns > 1 else None

        def resid(a, key):
            if ns > 1:
                return lambda i: ops.gemm_generic(a, layers[i % nl][key], gate=gate, rows_per_gate=S_LAT, split_out=parts)
            return lambda i: ops.gemm_generic(a, layers[i % nl][key], gate=gate, rows_per_gate=S_LAT, resid=x, out_f32=x)
        mine = dict(
            qkvg=lambda i: ops.gemm_sections(xn, layers[i % nl]["qkvg"],
                                             [(L.SEC_NORM_ROPE, qkvg[0], layers[i % nl]["qn"]), (L.SEC_NORM_ROPE, qkvg[1], layers[i % nl]["kn"]),
                                              (L.SEC_PLAIN, qkvg[2], None), (L.SEC_SIGMOID, qkvg[3], None)],
                                             head_dim=128, rope_heads=H // 2, rope=model.rope, rows_per_seq=S_LAT, eps=1e-5),
            wo=resid(xn, "wo"), w13=lambda i: ops.gemm_swiglu(xn, layers[i % nl]["w13"], hmid), w2=resid(hm, "w2"))
        shapes = dict(qkvg=(4 * D, D), wo=(D, D), w13=(2 * F, D), w2=(D, F))
        for name, (N, K) in shapes.items():
            a = hm if K == F else xn
            out = torch.empty((M, N), device="cuda", dtype=torch.bfloat16)
            us_mine = _event_us(mine[name], 2 * nl)
            us_cublas = _event_us(lambda i: torch.matmul(a, layers[i % nl][name].t(), out=out), 2 * nl)
            fl = 2.0 * M * N * K
            table[f"{name}_m{M}"] = dict(m=M, n=N, k=K, us=us_mine, tflops=fl / us_mine / 1e6, cublas_us=us_cublas,
                                         cublas_tflops=fl / us_cublas / 1e6, speed_vs_cublas=us_cublas / us_mine,
                                         split_k=(ns if name in ("wo", "w2") else 1))
    M, F2 = 3 * S_LAT, 2 * F
    a = torch.randn((M, D), device="cuda").to(torch.bfloat16)
    out = torch.empty((M, F), device="cuda", dtype=torch.bfloat16)
    n = 3 * nl
    sec = _event_us(lambda i: ops.gemm_swiglu(a, layers[i % nl]["w13"], out), n, warm=3) / 1e6
    flops = 2.0 * M * F2 * D
    ach = flops / sec / 1e12
    traffic, src = roofline_traffic()
    return dict(bound="tensor", kernel="gemm_pair_kernel<SWIGLU> (tcgen05 cta_group::2, 256x256 tiles) M=1920 N=11776 K=2048",
                achieved=ach, peak=pk["bf16"], unit="TFLOP/s", frac=ach / pk["bf16"], traffic=traffic,
                traffic_unit="bytes/launch (ncu dram__bytes_read.sum + dram__bytes_write.sum)", traffic_source=src,
                flop_per_launch=flops, peak_source=pk["src"] + " burst", launches_timed=n, us_per_launch=sec * 1e6,
                cublas_same_shape=table,
                cublas_note="torch.matmul(a, w.t()) bf16 = cuBLAS plain GEMM writing bf16 [M, N]; this library's figure "
                            "includes the fused epilogue (head RMSNorm + RoPE + sigmoid / SwiGLU / gated residual, split-K "
                            "partials where split_k > 1)")


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
                   "S1-DAC decode + D2H, " + str(dev[0].shape[0]) + " prompt(s); incremental_prefix: one CUDA graph per block (conditioning / prefix-KV "
                   "append + 40 steps) and per decode-stream step; recomputed_prefix = the reference's full prefix re-encode per "
                   "block, plain launches")
    return rec


def config3_batched(model, ae, pca, world, rank, local, total_prompts=64, per_forward=8):
    """BASELINE.json configs[2]: 64 independent prompts x 640 latents x 40 steps (distinct texts with valid lengths in
    [64, 512], distinct speaker references and seeds), dealt to the ranks in batches of 8 prompts per forward
    by serving.run_sharded -- data-parallel, no data-path collective.  Device-timed (CUDA events around this rank's
    batches, waveforms copied to the host inside the region), max over ranks."""
    from echo_tts_b200 import schema as S, serving
    from echo_tts_b200.inference import ae_decode, find_flattening_point
    from echo_tts_b200.sampler import sample_euler_cfg_independent_guidances as sample
    n_batches = total_prompts // per_forward
    lens = torch.randint(64, 513, (total_prompts,), generator=torch.Generator().manual_seed(64)).tolist()

    def make(bi):
        ids, mask, spk, smask = S.synth_request(per_forward, n_valid=N_TEXT_VALID, pad_to=L_TEXT, t_spk=T_SPK, seed=1000 + bi)
        for j in range(per_forward):
            n = lens[bi * per_forward + j]
            if n < N_TEXT_VALID:
                ids[j, n:] = 0; mask[j, n:] = False
            elif n > N_TEXT_VALID:
                extra = torch.randint(32, 127, (n - N_TEXT_VALID,), generator=torch.Generator().manual_seed(7000 + bi * 8 + j),
                                      dtype=torch.int32)
                ids[j, N_TEXT_VALID:n] = extra; mask[j, :n] = True
        return [t.pin_memory() for t in (ids, mask, spk, smask)]
    batches = [make(bi) for bi in range(n_batches)]

    def generate(item, bi):
        ids, mask, spk, smask = [t.cuda(non_blocking=True) for t in item]
        lat = sample(model, spk, smask, ids, mask, 500 + bi, **KW)
        wav = ae_decode(ae, pca, lat).cpu()
        return dict(batch=bi, samples=int(wav.shape[-1]) * per_forward,
                    crop=[int(find_flattening_point(lat[j])) for j in range(per_forward)])
    for i in range(3):                                                 # eager, capture, replay at the batch-8 shape
        generate(batches[0], 0)
    torch.cuda.synchronize()
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    res = serving.run_sharded(batches, generate, on_local_done=e1.record)      # the clock stops before the result gather
    torch.cuda.synchronize()
    ms = serving.max_over_ranks_ms(e0.elapsed_time(e1), torch.device("cuda", local))
    if rank != 0:
        return None
    audio = total_prompts * AUDIO_SEC_PER_PROMPT
    return dict(value=audio / (ms / 1e3), unit="audio-sec/sec", ms=ms, prompts=total_prompts, prompts_per_forward=per_forward,
                batches_per_rank=[len(serving.shard_indices(n_batches, r, world)) for r in range(world)],
                samples_out=sum(r["samples"] for r in res), flop=total_prompts * FLOP_GEN,
                tflops=total_prompts * FLOP_GEN / (ms / 1e3) / 1e12,
                what="BASELINE.json configs[2]: 64 prompts x 640 latents x 40 steps dealt round-robin in batches of 8 "
                     "(serving.run_sharded), no data-path collective; CUDA events around each rank's batches incl. the "
                     "waveform D2H, max over ranks")


def cfg_split_record(model, world, rank, local, reps=3):
    """Optional CFG-branch split (BASELINE configs[2] note): ranks 0..2 evaluate one guidance branch each and
    exchange the branch velocities once per guided step -- NCCL all-gather, or the fused NVLink peer-store exchange.  ms per
    clip (sampler only, configs[1] request) for both exchanges and for the single-GPU graph, max over the three ranks."""
    import torch.distributed as dist
    from echo_tts_b200 import serving
    from echo_tts_b200.sampler import sample_euler_cfg_independent_guidances as sample
    group = dist.new_group([0, 1, 2])
    if rank >= 3:
        return None
    ids, mask, spk, smask = [t.cuda() for t in synth_inputs(1)]
    noise = torch.randn(1, S_LAT, 80, generator=torch.Generator().manual_seed(3))
    dev = torch.device("cuda", local)

    def timed(fn):
        for _ in range(3):                                             # eager, capture, replay
            out = fn()
        torch.cuda.synchronize(); dist.barrier(group)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            out = fn()
        e1.record()
        torch.cuda.synchronize(); dist.barrier(group)
        return out, serving.max_over_ranks_ms(e0.elapsed_time(e1) / reps, dev, group)
    rec = {}
    outs = {}
    for ex in ("nccl", "p2p"):
        try:
            outs[ex], rec[f"ms_{ex}"] = timed(lambda: serving.sample_euler_cfg_split(model, spk, smask, ids, mask, 0, group=group,
                                                                                     exchange=ex, _noise=noise, **KW))
        except Exception as e:                                         # a latency option must never cost the bench line
            rec[f"error_{ex}"] = f"{type(e).__name__}: {e}"[:300]
    single, rec["ms_single_gpu_graph"] = timed(lambda: sample(model, spk, smask, ids, mask, 0, _noise=noise, **KW))
    if "nccl" in outs:
        gathered = [torch.empty_like(single) for _ in range(3)]
        dist.all_gather(gathered, outs["nccl"].contiguous(), group=group)
        rec["ranks_bit_identical"] = all(torch.equal(gathered[0], g) for g in gathered[1:])
        rec["rel_l2_vs_single_gpu"] = float((outs["nccl"] - single).norm() / single.norm())
        if "p2p" in outs:
            rec["p2p_equals_nccl"] = bool(torch.equal(outs["p2p"], outs["nccl"]))
    rec["what"] = ("serving.sample_euler_cfg_split on ranks 0-2: one branch per GPU on guided steps, one exchange of [1,640,80] "
                   "fp32 velocities per guided step (nccl = all-gather; p2p = tail kernel stores into peers' symmetric buffers), "
                   "combine + Euler redundantly on every rank; sampler only, ms per clip, max over the three ranks")
    return rec if rank == 0 else None


def run_gpu(args):
    from echo_tts_b200 import schema as S, _lib
    from echo_tts_b200.inference import (load_model_from_state_dict, load_fish_ae_from_state_dict, ae_decode, PCAState,
                                         find_flattening_point)
    from echo_tts_b200.sampler import sample_euler_cfg_independent_guidances as sample
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
    comp, mean, scale = S.synth_pca()
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
    c3 = config3_batched(model, ae, pca, world, rank, local) if (B == 1 and not args.no_config3) else None
    split = None
    if world >= 3 and B == 1 and not args.no_cfg_split:
        try:
            split = cfg_split_record(model, world, rank, local)
        except Exception as e:
            split = dict(error=f"{type(e).__name__}: {e}"[:300]) if rank == 0 else None
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
    if c3 is not None:
        line["config3"] = c3
    if split is not None:
        line["cfg_split"] = split
    if world == 1 and B == 1 and not args.no_eager_gpu:
        del model, ae
        torch.cuda.empty_cache()
        try:
            line["eager_torch_gpu"] = eager_gpu_port()
            line["eager_torch_gpu"]["speedup_of_this_engine"] = value / line["eager_torch_gpu"]["value"]
        except Exception as e:                                         # a baseline must never cost the bench line
            line["eager_torch_gpu"] = dict(error=f"{type(e).__name__}: {e}"[:300])
        model = ae = None
    if world == 1 and not args.no_cpu_baseline:      # rank 0 at N = 1 only (the contract)
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
    ap.add_argument("--no-config3", action="store_true", help="skip the BASELINE configs[2] record (64 prompts, 8 per forward)")
    ap.add_argument("--no-cfg-split", action="store_true", help="skip the CFG-branch split record (needs >= 3 GPUs)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    run_gpu(args)


if __name__ == "__main__":
    main()
