"""Euler sampler with independent text/speaker guidance on the sm_100a engine.

Host control loop of reference ``inference.py:446-560`` / ``inference_blockwise.py:14-123`` re-designed:
every t-only quantity (schedule, CFG window, AdaLN modulation table, speaker-KV multipliers) is computed
once per request; the three CFG branches share ONE conditioning cache (rows index it modulo B, masks drop a
modality); out_norm/out_proj/CFG-combine/rescale/Euler are one fused kernel; there is no host sync inside
the loop, and the whole 40-step loop replays as a single CUDA graph.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, Iterator, List, Optional, Sequence, Tuple

import torch

from . import _lib as L
from . import ops
from .model import EchoDiT, KVCache, KVList, BF16, F32

INIT_SCALE = 0.999  # inference.py:470


class StepPlan:
    """Host-side schedule logic, bit-exact with the reference's fp32 device-scalar comparisons."""

    def __init__(self, device, num_steps: int, cfg_min_t: float, cfg_max_t: float, speaker_kv_scale, speaker_kv_max_layers,
                 speaker_kv_min_t, num_layers: int, initial_multiplier: Optional[List[float]] = None):
        t_sched = torch.linspace(1.0, 0.0, num_steps + 1, device=device) * INIT_SCALE      # inference.py:477
        th = t_sched.cpu()
        t, tn = th[:-1], th[1:]
        self.t_dev = t_sched[:-1].contiguous()
        self.has_cfg: List[bool] = ((t >= cfg_min_t) * (t <= cfg_max_t)).tolist()          # inference.py:511
        sched = torch.zeros((num_steps, 4), device=device, dtype=F32)
        sched[:, 0] = t_sched[:-1]
        sched[:, 1] = t_sched[1:] - t_sched[:-1]                                           # inference.py:558
        self.sched = sched
        n_scaled = num_layers if speaker_kv_max_layers is None else min(speaker_kv_max_layers, num_layers)
        mult = list(initial_multiplier) if initial_multiplier is not None else [1.0] * num_layers
        self.kv_mult_rows: List[List[float]] = []
        self.unscale_after: Optional[int] = None
        if speaker_kv_scale is not None:
            for l in range(n_scaled):                                                      # inference.py:485-486
                mult[l] *= speaker_kv_scale
            crossing = ((tn < speaker_kv_min_t) * (t >= speaker_kv_min_t)).tolist()        # inference.py:546-550
        else:
            crossing = [False] * num_steps
        for i in range(num_steps):
            self.kv_mult_rows.append(list(mult))
            if crossing[i]:
                if self.unscale_after is None:
                    self.unscale_after = i
                for l in range(n_scaled):
                    mult[l] *= 1.0 / speaker_kv_scale                                      # inference.py:551-553
        self.final_multiplier = mult
        self.uses_kv_scale = speaker_kv_scale is not None
        self.kv_mult = torch.tensor(self.kv_mult_rows, device=device, dtype=F32) if self.uses_kv_scale else None


def _prep_masks(model: EchoDiT, text_mask: torch.Tensor, speaker_mask: torch.Tensor):
    """Row layout (cond | text-dropped | speaker-dropped), inference.py:494-497; speaker mask at patch rate
    (model.py:581).  Returns uint8 masks + per-row extents for 3B rows; the first B rows serve plain steps."""
    dev = model.device
    p = model.cfg.speaker_patch_size
    tm = text_mask.to(dev).to(torch.uint8).contiguous()
    sm = speaker_mask.to(dev)[..., ::p].to(torch.uint8).contiguous()
    tm3 = torch.cat([tm, torch.zeros_like(tm), tm], 0).contiguous()
    sm3 = torch.cat([sm, sm, torch.zeros_like(sm)], 0).contiguous()
    return tm3, ops.mask_extent(tm3), sm3, ops.mask_extent(sm3)


def run_steps(model: EchoDiT, x_t: torch.Tensor, plan: StepPlan, mod: torch.Tensor, B: int, masks, kv_text: KVCache,
              kv_spk: KVCache, p: dict, *, start_pos: int = 0, kv_latent: Optional[KVCache] = None, n_prefix: int = 0,
              trace: Optional[list] = None, teacher: Optional[Dict[int, torch.Tensor]] = None,
              guidance: Optional[torch.Tensor] = None) -> list:
    """The hot loop (inference.py:508-558): x_t fp32 [B, S, latent] updated in place; launches only.  Without a trace
    the whole loop is ONE native call (``echo_dit_sample``: per step the DiT forward over 3B or B rows + the fused
    out-projection / CFG combine / rescale / Euler tail); with a trace, one ``echo_dit_step`` per step, and
    ``teacher[i]`` (parity tests) replaces x_t before step i -- velocities at the reference's own x_t.
    ``guidance``: device fp32 [4] (cfg scales, rescale k / sigma) read by the tail kernel instead of the immediates, so
    a captured graph serves every guidance setting.  Returns the objects the launches point into (workspaces, pointer
    arrays): a CUDA graph captured around this call must keep them alive."""
    cfg = model.cfg
    tm3, te3, sm3, se3 = masks
    S = x_t.shape[1]
    use_rescale = p["rescale_k"] is not None and p["rescale_sigma"] is not None
    n = len(plan.has_cfg)
    lib = L.load()
    torch.cuda.nvtx.range_push(f"echo.euler_steps[{n}]")

    def step(R: int, i: int, **tail_kw):
        tail = ops.tail_desc(model.out_norm, model.out_w, model.out_b, positions=B * S, branches=R // B,
                             eps=cfg.norm_eps, x_t=x_t, sched=plan.sched[i:], cfg_text=p["cfg_scale_text"],
                             cfg_speaker=p["cfg_scale_speaker"], rescale_k=p["rescale_k"] if use_rescale else None,
                             rescale_sigma=p["rescale_sigma"] if use_rescale else None, guidance=guidance, **tail_kw)
        return model.step_desc(x_t, mod[i:i + 1], B, R, text_mask=tm3, text_extent=te3, spk_mask=sm3, spk_extent=se3,
                               kv_text=kv_text, kv_speaker=kv_spk, kv_latent=kv_latent, start_pos=start_pos,
                               n_prefix=n_prefix, spk_scale=(plan.kv_mult[i] if plan.uses_kv_scale else None),
                               tail=tail)

    if trace is None:
        assert mod.shape[0] >= n and plan.sched.is_contiguous()
        d3, keep3 = step(3 * B, 0) if any(plan.has_cfg) else (None, None)
        d1, keep1 = step(B, 0) if not all(plan.has_cfg) else (None, None)
        flags = (C.c_uint8 * max(n, 1))(*[int(f) for f in plan.has_cfg])
        L.check(lib.echo_dit_sample(C.byref(d3) if d3 is not None else None, C.byref(d1) if d1 is not None else None,
                                    n, C.cast(flags, C.c_void_p), mod.stride(0),
                                    plan.kv_mult.stride(0) if plan.uses_kv_scale else 0, plan.sched.stride(0),
                                    L.stream()))
        torch.cuda.nvtx.range_pop()
        return [keep3, keep1]
    for i, has_cfg in enumerate(plan.has_cfg):
        R = 3 * B if has_cfg else B
        v_out = torch.empty((R * S, cfg.latent_size), device=x_t.device, dtype=F32)
        v_comb = torch.empty((B * S, cfg.latent_size), device=x_t.device, dtype=F32)
        if teacher is not None and i in teacher:
            x_t.copy_(teacher[i].to(x_t.device, F32).view_as(x_t))
        x_in = x_t.clone()
        d, keep = step(R, i, v_out=v_out, v_comb_out=v_comb)
        L.check(lib.echo_dit_step(C.byref(d), L.stream()))
        trace.append(dict(i=i, has_cfg=has_cfg, x_in=x_in, v_branches=v_out.view(R, S, -1), v=v_comb.view(B, S, -1)))
    torch.cuda.nvtx.range_pop()
    return []


class _GraphEntry:
    def __init__(self):
        self.calls = 0
        self.graph: Optional[torch.cuda.CUDAGraph] = None
        self.static: Dict[str, torch.Tensor] = {}


def graph_cache_capacity() -> int:
    """Captured sampler graphs kept per model (least recently used evicted).  An entry's private pool pins the 24-layer
    text + speaker KV caches and the encoder scratch of its shape (~160 MB at 53 speaker patches, > 450 MB at 1600)."""
    return max(1, int(os.environ.get("ECHO_GRAPH_CACHE", "8")))


def _graph_entry(model: EchoDiT, key) -> _GraphEntry:
    """LRU lookup in the per-model graph cache.  Keys hold SHAPES and the schedule only (text / speaker lengths as the
    caller bucketed them -- audio.request_buckets -- steps, the CFG window's step set, rescale on/off, kv-scale on/off);
    guidance scales, rescale constants, speaker-KV multipliers and the t schedule are device data refreshed per call."""
    from collections import OrderedDict
    cache = model.__dict__.get("_graphs")
    if not isinstance(cache, OrderedDict):
        cache = model.__dict__["_graphs"] = OrderedDict(cache or {})
    ent = cache.get(key)
    if ent is None:
        ent = cache[key] = _GraphEntry()
    cache.move_to_end(key)
    while len(cache) > graph_cache_capacity():
        _, old = cache.popitem(last=False)
        old.graph = None            # drops the graph, its private memory pool and the static input buffers
        old.static.clear()
    return ent


def _use_graphs() -> bool:
    return os.environ.get("ECHO_CUDA_GRAPH", "1") != "0"


@torch.inference_mode()
def sample_euler_cfg_independent_guidances(
        model: EchoDiT, speaker_latent: torch.Tensor, speaker_mask: torch.Tensor, text_input_ids: torch.Tensor,
        text_mask: torch.Tensor, rng_seed: int, num_steps: int, cfg_scale_text: float, cfg_scale_speaker: float,
        cfg_min_t: float, cfg_max_t: float, truncation_factor: Optional[float], rescale_k: Optional[float],
        rescale_sigma: Optional[float], speaker_kv_scale: Optional[float], speaker_kv_max_layers: Optional[int],
        speaker_kv_min_t: Optional[float], sequence_length: Optional[int] = None, *,
        _noise: Optional[torch.Tensor] = None, _trace: Optional[list] = None,
        _teacher: Optional[Dict[int, torch.Tensor]] = None) -> torch.Tensor:
    """Drop-in for reference ``inference.py:446-560`` (same positional/keyword arguments, fp32 [B, S, 80] out).
    ``_noise`` injects x_T, ``_trace`` collects per-step velocities and ``_teacher`` (with a trace) forces x_t at the
    given steps (parity tests only)."""
    if sequence_length is None:
        sequence_length = 640
    if speaker_kv_scale is not None and speaker_kv_min_t is None:
        raise TypeError("'<' not supported between instances of 'Tensor' and 'NoneType'")   # inference.py:548
    cfg, dev = model.cfg, model.device
    B, S = text_input_ids.shape[0], sequence_length
    p = dict(cfg_scale_text=float(cfg_scale_text), cfg_scale_speaker=float(cfg_scale_speaker), rescale_k=rescale_k,
             rescale_sigma=rescale_sigma)
    plan = StepPlan(dev, num_steps, cfg_min_t, cfg_max_t, speaker_kv_scale, speaker_kv_max_layers, speaker_kv_min_t,
                    cfg.num_layers)
    if _noise is not None:
        x_t = _noise.to(dev, F32).clone()
    else:
        rng = torch.Generator(device=dev).manual_seed(rng_seed)                            # inference.py:475
        x_t = torch.randn((B, S, cfg.latent_size), device=dev, dtype=F32, generator=rng)   # inference.py:499
    if truncation_factor is not None:
        x_t = x_t * truncation_factor

    def conditioning(ids, tmask, spk, smask):
        torch.cuda.nvtx.range_push("echo.conditioning")
        kv_t, kv_s = model.kv_caches_text_speaker(ids, tmask, spk)                         # inference.py:482-483
        out = kv_t, kv_s, _prep_masks(model, tmask, smask)
        torch.cuda.nvtx.range_pop()
        return out

    use_rescale = rescale_k is not None and rescale_sigma is not None
    key = ("euler", B, S, text_input_ids.shape[1], speaker_latent.shape[1], num_steps, tuple(plan.has_cfg), use_rescale,
           plan.uses_kv_scale)
    ent = _graph_entry(model, key)
    ent.calls += 1
    if _trace is not None or not _use_graphs() or ent.calls == 1:
        # plain launches: parity traces, ECHO_CUDA_GRAPH=0, and the first request with a new shape (warms kernels up)
        if _trace is not None or not _use_graphs():
            ent.calls -= 1
        kv_text, kv_spk, masks = conditioning(text_input_ids, text_mask, speaker_latent, speaker_mask)
        mod = model.build_mod_table(plan.t_dev)
        run_steps(model, x_t, plan, mod, B, masks, kv_text, kv_spk, p, trace=_trace, teacher=_teacher)
        return x_t.view(B, S, cfg.latent_size)

    # ---- CUDA-graph path: conditioning encoders + all Euler steps are ONE captured graph over static input buffers,
    #      keyed by everything baked into the captured launches (shapes, schedule, guidance constants) ----
    st = ent.static
    if ent.graph is None:
        st["ids"] = text_input_ids.to(dev, torch.int32).contiguous().clone()
        st["tmask"] = text_mask.to(dev).contiguous().clone()
        st["spk"] = speaker_latent.to(dev, F32).contiguous().clone()
        st["smask"] = speaker_mask.to(dev).contiguous().clone()
        st["x_t"] = x_t.clone()
        st["mod"] = model.build_mod_table(plan.t_dev)     # depends on the schedule only (part of the key)
        st["plan"] = plan
        st["guidance"] = torch.zeros(4, device=dev, dtype=F32)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            kv_text, kv_spk, masks = conditioning(st["ids"], st["tmask"], st["spk"], st["smask"])
            st["keep"] = run_steps(model, st["x_t"], plan, st["mod"], B, masks, kv_text, kv_spk, p,
                                   guidance=st["guidance"])
        ent.graph = g
    st["ids"].copy_(text_input_ids)
    st["tmask"].copy_(text_mask)
    st["spk"].copy_(speaker_latent)
    st["smask"].copy_(speaker_mask)
    st["x_t"].copy_(x_t)
    st["guidance"].copy_(torch.tensor([p["cfg_scale_text"], p["cfg_scale_speaker"], rescale_k if use_rescale else 1.0,
                                       rescale_sigma if use_rescale else 0.0], dtype=F32), non_blocking=True)
    if plan.uses_kv_scale:
        st["plan"].kv_mult.copy_(plan.kv_mult)
    st["plan"].sched.copy_(plan.sched)
    torch.cuda.nvtx.range_push("echo.sampler_graph_replay")
    ent.graph.replay()
    torch.cuda.nvtx.range_pop()
    return st["x_t"].clone().view(B, S, cfg.latent_size)


def _blockwise_graphed(model: EchoDiT, ent: "_GraphEntry", speaker_latent, speaker_mask, text_input_ids, text_mask, rng,
                       block_sizes, num_steps, cfg_min_t, cfg_max_t, truncation_factor, speaker_kv_scale,
                       speaker_kv_max_layers, speaker_kv_min_t, p) -> Iterator[Tuple[int, int, torch.Tensor]]:
    """The block loop of ``blockwise_blocks`` as one CUDA graph per block (captured at the second request of a key, replayed
    afterwards).  Yields the STATIC prefix buffer: consume ``prefix[:, start:stop]`` before asking for the next request."""
    from .streaming import PrefixKVStream
    cfg, dev = model.cfg, model.device
    B = text_input_ids.shape[0]
    use_rescale = p["rescale_k"] is not None and p["rescale_sigma"] is not None
    st = ent.static
    if "blocks" not in st:
        st["ids"] = text_input_ids.to(dev, torch.int32).contiguous().clone()
        st["tmask"] = text_mask.to(dev).contiguous().clone()
        st["spk"] = speaker_latent.to(dev, F32).contiguous().clone()
        st["smask"] = speaker_mask.to(dev).contiguous().clone()
        st["guidance"] = torch.zeros(4, device=dev, dtype=F32)
        st["prefix"] = torch.zeros((B, sum(block_sizes), cfg.latent_size), device=dev, dtype=F32)
        st["pstream"] = PrefixKVStream(model, B, sum(block_sizes))
        st["mod"] = None
        st["blocks"] = []
    st["ids"].copy_(text_input_ids); st["tmask"].copy_(text_mask)
    st["spk"].copy_(speaker_latent); st["smask"].copy_(speaker_mask)
    st["guidance"].copy_(torch.tensor([p["cfg_scale_text"], p["cfg_scale_speaker"], p["rescale_k"] if use_rescale else 1.0,
                                       p["rescale_sigma"] if use_rescale else 0.0], dtype=F32), non_blocking=True)
    mult, start = None, 0
    for bi, bs in enumerate(block_sizes):
        plan = StepPlan(dev, num_steps, cfg_min_t, cfg_max_t, speaker_kv_scale, speaker_kv_max_layers, speaker_kv_min_t,
                        cfg.num_layers, initial_multiplier=mult)                     # inference_blockwise.py:68-70
        mult = plan.final_multiplier
        x_t = torch.randn((B, bs, cfg.latent_size), device=dev, dtype=F32, generator=rng)     # inference_blockwise.py:76
        if truncation_factor is not None:
            x_t = x_t * truncation_factor
        if bi >= len(st["blocks"]):
            if st["mod"] is None:
                st["mod"] = model.build_mod_table(plan.t_dev)
            blk = dict(x_t=x_t.clone(), plan=plan)
            pstream = st["pstream"]
            if bi == 0:
                pstream.patches = 0
                pstream.pending = torch.empty((B, 0, cfg.latent_size), device=dev, dtype=F32)
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                if bi == 0:
                    st["cond"] = (*model.kv_caches_text_speaker(st["ids"], st["tmask"], st["spk"]),
                                  _prep_masks(model, st["tmask"], st["smask"]))
                else:
                    pstream.append(st["blocks"][bi - 1]["x_t"])
                kv_text, kv_spk, masks = st["cond"]
                kv_lat, n_prefix = (pstream.kv, pstream.patches) if pstream.patches > 0 else (None, 0)
                blk["keep"] = run_steps(model, blk["x_t"], plan, st["mod"], B, masks, kv_text, kv_spk, p, start_pos=start,
                                        kv_latent=kv_lat, n_prefix=n_prefix, guidance=st["guidance"])
                st["prefix"][:, start:start + bs].copy_(blk["x_t"])
            blk["graph"] = g
            st["blocks"].append(blk)
        blk = st["blocks"][bi]
        blk["x_t"].copy_(x_t)
        blk["plan"].sched.copy_(plan.sched)
        if plan.uses_kv_scale:
            blk["plan"].kv_mult.copy_(plan.kv_mult)
        blk["graph"].replay()
        yield start, start + bs, st["prefix"]
        start += bs


@torch.inference_mode()
def blockwise_blocks(
        model: EchoDiT, speaker_latent: torch.Tensor, speaker_mask: torch.Tensor, text_input_ids: torch.Tensor,
        text_mask: torch.Tensor, rng_seed: int, block_sizes: Sequence[int], num_steps: int, cfg_scale_text: float,
        cfg_scale_speaker: float, cfg_min_t: float, cfg_max_t: float, truncation_factor: Optional[float],
        rescale_k: Optional[float], rescale_sigma: Optional[float], speaker_kv_scale: Optional[float],
        speaker_kv_max_layers: Optional[int], speaker_kv_min_t: Optional[float],
        continuation_latent: Optional[torch.Tensor] = None, *, _noises: Optional[Sequence[torch.Tensor]] = None,
        incremental_prefix: Optional[bool] = None, _trace: Optional[list] = None) -> Iterator[Tuple[int, int, torch.Tensor]]:
    """The block loop of reference ``inference_blockwise.py:14-123`` as a generator: yields ``(start, stop, prefix)``
    after each block (``prefix[:, :stop]`` is final), which is what a streaming caller decodes.  Differences in HOW,
    not WHAT: the latent-prefix KV is encoded for B rows (the reference encodes three identical copies, :72-74),
    un-generated patches are skipped by length instead of masked, and -- when every block boundary falls on a patch
    boundary -- the prefix KV is grown incrementally (``streaming.PrefixKVStream``: the encoder is causal, so only the
    new patches are encoded) instead of being recomputed in full at every block."""
    if speaker_kv_scale is not None and speaker_kv_min_t is None:
        raise TypeError("'<' not supported between instances of 'Tensor' and 'NoneType'")
    cfg, dev = model.cfg, model.device
    B = text_input_ids.shape[0]
    p = dict(cfg_scale_text=float(cfg_scale_text), cfg_scale_speaker=float(cfg_scale_speaker), rescale_k=rescale_k,
             rescale_sigma=rescale_sigma)
    rng = torch.Generator(device=dev).manual_seed(rng_seed)
    prefix = torch.zeros((B, sum(block_sizes), cfg.latent_size), device=dev, dtype=F32)
    start = 0
    if continuation_latent is not None:
        prefix = torch.cat([continuation_latent.to(dev, F32), prefix], 1)
        start = continuation_latent.shape[1]
    ps = cfg.speaker_patch_size
    aligned = start % ps == 0 and all(bs % ps == 0 for bs in block_sizes)
    if incremental_prefix is None:
        incremental_prefix = os.environ.get("ECHO_PREFIX_INCREMENTAL", "1") != "0"
    # ---- CUDA-graph path (the serving case: no continuation, patch-aligned blocks, incremental prefix): one captured graph
    #      per block -- block 0 = conditioning encoders + its Euler steps, block i = append block i-1 to the prefix KV + its
    #      Euler steps -- over static buffers, keyed by shapes and schedule only.  First request of a key: plain launches.
    if (_use_graphs() and _noises is None and _trace is None and start == 0 and aligned and incremental_prefix
            and model.has_blockwise and len(block_sizes) > 0 and min(block_sizes) >= ps):
        use_rescale = rescale_k is not None and rescale_sigma is not None
        plan0 = StepPlan(dev, num_steps, cfg_min_t, cfg_max_t, speaker_kv_scale, speaker_kv_max_layers, speaker_kv_min_t,
                         cfg.num_layers)
        key = ("blockwise", B, tuple(int(b) for b in block_sizes), text_input_ids.shape[1], speaker_latent.shape[1],
               num_steps, tuple(plan0.has_cfg), use_rescale, plan0.uses_kv_scale)
        ent = _graph_entry(model, key)
        ent.calls += 1
        if ent.calls >= 2:
            yield from _blockwise_graphed(model, ent, speaker_latent, speaker_mask, text_input_ids, text_mask, rng, block_sizes,
                                          num_steps, cfg_min_t, cfg_max_t, truncation_factor, speaker_kv_scale,
                                          speaker_kv_max_layers, speaker_kv_min_t, p)
            return
    kv_text, kv_spk = model.kv_caches_text_speaker(text_input_ids, text_mask, speaker_latent)
    masks = _prep_masks(model, text_mask, speaker_mask)
    pstream = None
    if incremental_prefix and aligned and model.has_blockwise and prefix.shape[1] >= ps:
        from .streaming import PrefixKVStream
        pstream = PrefixKVStream(model, B, prefix.shape[1])
        if start > 0:
            pstream.append(prefix[:, :start])
    mult = None
    mod = None
    for bi, bs in enumerate(block_sizes):
        # the reference re-applies speaker_kv_scale at every block (inference_blockwise.py:68-70): carry the multiplier
        plan = StepPlan(dev, num_steps, cfg_min_t, cfg_max_t, speaker_kv_scale, speaker_kv_max_layers, speaker_kv_min_t,
                        cfg.num_layers, initial_multiplier=mult)
        mult = plan.final_multiplier
        if mod is None:
            mod = model.build_mod_table(plan.t_dev)
        if pstream is not None:
            kv_lat, n_prefix = (pstream.kv, pstream.patches) if pstream.patches > 0 else (None, 0)
        else:
            kv_lat = model.get_kv_cache_latent(prefix) if prefix.shape[1] >= ps else None
            n_prefix = 0
            if kv_lat is not None:
                n_prefix = min(kv_lat[0][0].shape[1], -(-start // ps))                         # model.py:243-244
        if _noises is not None:
            x_t = _noises[bi].to(dev, F32).clone()
        else:
            x_t = torch.randn((B, bs, cfg.latent_size), device=dev, dtype=F32, generator=rng)
        if truncation_factor is not None:
            x_t = x_t * truncation_factor
        run_steps(model, x_t, plan, mod, B, masks, kv_text, kv_spk, p, start_pos=start, kv_latent=kv_lat,
                  n_prefix=n_prefix, trace=_trace)
        prefix[:, start:start + bs] = x_t
        if pstream is not None and bi + 1 < len(block_sizes):
            pstream.append(x_t)
        yield start, start + bs, prefix
        start += bs


@torch.inference_mode()
def sample_blockwise_euler_cfg_independent_guidances(
        model: EchoDiT, speaker_latent: torch.Tensor, speaker_mask: torch.Tensor, text_input_ids: torch.Tensor,
        text_mask: torch.Tensor, rng_seed: int, block_sizes: Sequence[int], num_steps: int, cfg_scale_text: float,
        cfg_scale_speaker: float, cfg_min_t: float, cfg_max_t: float, truncation_factor: Optional[float],
        rescale_k: Optional[float], rescale_sigma: Optional[float], speaker_kv_scale: Optional[float],
        speaker_kv_max_layers: Optional[int], speaker_kv_min_t: Optional[float],
        continuation_latent: Optional[torch.Tensor] = None, *, _noises: Optional[Sequence[torch.Tensor]] = None,
        on_block=None, incremental_prefix: Optional[bool] = None, _trace: Optional[list] = None) -> torch.Tensor:
    """Drop-in for reference ``inference_blockwise.py:14-123`` (see ``blockwise_blocks`` for the loop);
    ``on_block(start, stop, prefix)`` (optional) is the streaming hook."""
    if speaker_kv_scale is not None and speaker_kv_min_t is None:
        raise TypeError("'<' not supported between instances of 'Tensor' and 'NoneType'")
    prefix = None
    for start, stop, prefix in blockwise_blocks(
            model, speaker_latent, speaker_mask, text_input_ids, text_mask, rng_seed, block_sizes, num_steps,
            cfg_scale_text, cfg_scale_speaker, cfg_min_t, cfg_max_t, truncation_factor, rescale_k, rescale_sigma,
            speaker_kv_scale, speaker_kv_max_layers, speaker_kv_min_t, continuation_latent, _noises=_noises,
            incremental_prefix=incremental_prefix, _trace=_trace):
        if on_block is not None:
            on_block(start, stop, prefix)
    if prefix is not None:
        prefix = prefix.clone()      # the graph path yields its static prefix buffer
    if prefix is None:          # no blocks: the (possibly empty) continuation comes back unchanged
        cfg, dev = model.cfg, model.device
        B = text_input_ids.shape[0]
        prefix = torch.zeros((B, 0, cfg.latent_size), device=dev, dtype=F32)
        if continuation_latent is not None:
            prefix = continuation_latent.to(dev, F32).clone()
    return prefix
