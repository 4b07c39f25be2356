"""Multi-GPU serving: independent prompts, data-parallel, one process per GPU.

The sampling path has no cross-prompt operation (every tensor carries a leading batch axis: reference
``model.py:215-268``, ``inference.py:473-558``), so prompts are dealt round-robin to ranks, weights are replicated and
there is NO collective on the data path; only the finished waveforms (host tensors of different lengths) are gathered
on rank 0.  ``torch.distributed`` is plumbing here: NCCL on the GPU box, gloo in the CPU tests.
"""
from __future__ import annotations

from typing import Callable, List, Optional, Sequence, TypeVar

import torch
import torch.distributed as dist

T = TypeVar("T")
R = TypeVar("R")


def shard_indices(n_items: int, rank: int, world: int) -> List[int]:
    """Round-robin deal: item i belongs to rank i % world (keeps per-rank loads within one item of each other)."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world of {world}")
    return list(range(rank, n_items, world))


def run_sharded(items: Sequence[T], generate: Callable[[T, int], R], *, group=None, dst: int = 0,
                on_local_done: Optional[Callable[[], None]] = None) -> Optional[List[R]]:
    """Run ``generate(item, global_index)`` for this rank's share of ``items`` and gather the results, in the
    original order, on rank ``dst`` (other ranks get None).  Results must be picklable host objects (e.g. CPU
    waveforms): gathering is control-plane traffic, not part of the timed data path -- ``on_local_done`` (optional) is
    called when this rank's share is finished and before the gather, which is where a benchmark stops its clock."""
    if not (dist.is_available() and dist.is_initialized()):
        out = [generate(it, i) for i, it in enumerate(items)]
        if on_local_done is not None:
            on_local_done()
        return out
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    mine = shard_indices(len(items), rank, world)
    local = [(i, generate(items[i], i)) for i in mine]
    if on_local_done is not None:
        on_local_done()
    gathered: Optional[list] = [None] * world if rank == dst else None
    dist.gather_object(local, gathered, dst=dst, group=group)
    if rank != dst:
        return None
    out: List[Optional[R]] = [None] * len(items)
    for part in gathered:
        for i, r in part:
            out[i] = r
    assert all(r is not None for r in out) or len(items) == 0
    return out  # type: ignore[return-value]


def max_over_ranks_ms(ms: float, device: Optional[torch.device] = None, group=None) -> float:
    """Device-timed milliseconds -> max over ranks (the number every multi-GPU figure is quoted on)."""
    if not (dist.is_available() and dist.is_initialized()):
        return ms
    t = torch.tensor([ms], dtype=torch.float64, device=device or "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return float(t.item())


# ----------------------------------------------------------------------------------------------------------------
# Optional latency mode: CFG-branch split over three GPUs.
# ----------------------------------------------------------------------------------------------------------------
_P2P_CACHE: dict = {}


def _p2p_buffers(group, rank: int, P: int, latent: int, dev):
    """Symmetric (peer-mapped) gather buffers of the fused exchange, one rendezvous per (group, shape)."""
    import torch.distributed._symmetric_memory as symm
    key = (id(group), P, latent)
    ent = _P2P_CACHE.get(key)
    if ent is None:
        F32 = torch.float32
        v3_all = symm.empty((2, 3, P, latent), dtype=F32, device=dev)            # double-buffered by guided-step parity
        flags = symm.empty((4,), dtype=torch.int32, device=dev)                  # counter s: CTAs of sender s seen so far
        hv, hf = symm.rendezvous(v3_all, group), symm.rendezvous(flags, group)
        others = [r for r in range(3) if r != rank]
        peer_v = [hv.get_buffer(r, (2, 3, P, latent), F32) for r in others]
        peer_f = [hf.get_buffer(r, (4,), torch.int32) for r in others]
        ent = (v3_all, flags, hf, peer_v, peer_f, hv)
        _P2P_CACHE[key] = ent
    return ent[:5]


def branch_rows(rank: int, B: int) -> slice:
    """Rows of the (cond | text-dropped | speaker-dropped) mask layout (reference ``inference.py:494-497``) that branch
    rank ``rank`` (0, 1, 2) evaluates."""
    if rank not in (0, 1, 2):
        raise ValueError("the CFG-branch split uses exactly three ranks")
    return slice(rank * B, (rank + 1) * B)


@torch.inference_mode()
def sample_euler_cfg_split(model, speaker_latent, speaker_mask, text_input_ids, text_mask, rng_seed: int, num_steps: int,
                           cfg_scale_text: float, cfg_scale_speaker: float, cfg_min_t: float, cfg_max_t: float,
                           truncation_factor, rescale_k, rescale_sigma, speaker_kv_scale, speaker_kv_max_layers,
                           speaker_kv_min_t, sequence_length: Optional[int] = None, *, group=None,
                           exchange: str = "nccl", use_graph: bool = True,
                           _noise: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``sample_euler_cfg_independent_guidances`` with the three guidance branches of every CFG step evaluated on
    three GPUs (ranks 0/1/2 of ``group`` = cond / text-dropped / speaker-dropped; every rank holds the full model and
    receives the same inputs and seed).  Per CFG step: one forward over B rows instead of 3B, ONE all-gather of the
    branch velocities ([B, S, 80] fp32 per rank) over NCCL/NVLink, then the combine + rescale + Euler update
    redundantly on every rank, so x_t stays bit-identical across ranks without a broadcast.  Un-guided steps run
    replicated.  ``exchange="p2p"`` fuses the exchange into the kernels instead of calling NCCL: the tail kernel of
    every rank stores its velocities straight into the other two ranks' gather buffers (symmetric memory, NVLink peer
    stores) and bumps a counter there; the combine kernel waits on its two counters.  No collective launch at all.  A latency feature for few prompts (about 3x fewer tokens per GPU on guided steps); for throughput
    use ``run_sharded`` instead."""
    from . import ops
    from .sampler import StepPlan, _prep_masks, INIT_SCALE  # noqa: F401
    F32 = torch.float32
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    if world != 3:
        raise ValueError(f"CFG-branch split needs a process group of exactly 3 ranks, got {world}")
    if sequence_length is None:
        sequence_length = 640
    cfg, dev = model.cfg, model.device
    B, S = text_input_ids.shape[0], sequence_length
    plan = StepPlan(dev, num_steps, cfg_min_t, cfg_max_t, speaker_kv_scale, speaker_kv_max_layers, speaker_kv_min_t,
                    cfg.num_layers)
    if _noise is not None:
        x_t = _noise.to(dev, F32).clone()
    else:
        rng = torch.Generator(device=dev).manual_seed(rng_seed)
        x_t = torch.randn((B, S, cfg.latent_size), device=dev, dtype=F32, generator=rng)
    if truncation_factor is not None:
        x_t = x_t * truncation_factor
    use_rescale = rescale_k is not None and rescale_sigma is not None
    mine = branch_rows(rank, B)
    P = B * S
    if exchange == "p2p":
        lib = __import__("echo_tts_b200._lib", fromlist=["load"]).load()
        v3_all, flags, hf, peer_v, peer_f = _p2p_buffers(group, rank, P, cfg.latent_size, dev)
        others = [r for r in range(3) if r != rank]
        ctas = int(lib.echo_cfg_tail_ctas(P))
    elif exchange == "nccl":
        v3 = torch.empty((3, P, cfg.latent_size), device=dev, dtype=F32)
    else:
        raise ValueError(f"exchange must be 'nccl' or 'p2p', got {exchange!r}")

    def body(ids, tmask, spk, smask, x):
        """conditioning + the whole Euler loop on x (in place): launches only (CUDA-graph capturable with p2p)"""
        kv_text = model.get_kv_cache_text(ids, tmask)
        kv_spk = model.get_kv_cache_speaker(spk)
        tm3, te3, sm3, se3 = _prep_masks(model, tmask, smask)
        n_guided = 0
        for i, has_cfg in enumerate(plan.has_cfg):
            rows = mine if has_cfg else slice(0, B)
            h, pending = model.forward_hidden(x, mod[i:i + 1], B, text_mask=tm3[rows], text_extent=te3[rows],
                                              spk_mask=sm3[rows], spk_extent=se3[rows], kv_text=kv_text, kv_speaker=kv_spk,
                                              spk_scale=(plan.kv_mult[i] if plan.uses_kv_scale else None))
            if has_cfg and exchange == "p2p":
                par, n_guided = n_guided & 1, n_guided + 1
                ops.cfg_euler_tail(h, model.out_norm, model.out_w, model.out_b, positions=P, branches=1, eps=cfg.norm_eps,
                                   v_out=v3_all[par, rank], update_x=False, delta=pending,
                                   push_to=[(pv[par, rank], pf[rank:rank + 1]) for pv, pf in zip(peer_v, peer_f)])
                ops.cfg_combine_euler(v3_all[par], x.view(P, -1), plan.sched[i], cfg_text=float(cfg_scale_text),
                                      cfg_speaker=float(cfg_scale_speaker), rescale_k=rescale_k if use_rescale else None,
                                      rescale_sigma=rescale_sigma if use_rescale else None,
                                      wait=(flags[others[0]:others[0] + 1], flags[others[1]:others[1] + 1], ctas * n_guided))
            elif has_cfg:
                ops.cfg_euler_tail(h, model.out_norm, model.out_w, model.out_b, positions=P, branches=1, eps=cfg.norm_eps,
                                   v_out=v3[rank], update_x=False, delta=pending)
                dist.all_gather_into_tensor(v3.view(-1), v3[rank].view(-1), group=group)
                ops.cfg_combine_euler(v3, x.view(P, -1), plan.sched[i], cfg_text=float(cfg_scale_text),
                                      cfg_speaker=float(cfg_scale_speaker), rescale_k=rescale_k if use_rescale else None,
                                      rescale_sigma=rescale_sigma if use_rescale else None)
            else:
                ops.cfg_euler_tail(h, model.out_norm, model.out_w, model.out_b, positions=B * S, branches=1, eps=cfg.norm_eps,
                                   x_t=x, sched=plan.sched[i], rescale_k=rescale_k if use_rescale else None,
                                   rescale_sigma=rescale_sigma if use_rescale else None, delta=pending)

    if exchange != "p2p" or not use_graph:
        mod = model.build_mod_table(plan.t_dev)
        if exchange == "p2p":
            flags.zero_()
            torch.cuda.synchronize()
            hf.barrier()                                                         # nobody pushes before everyone zeroed
        body(text_input_ids, text_mask, speaker_latent, speaker_mask, x_t)
    else:
        # fused exchange: no collective call inside the loop, so conditioning + all 40 steps replay as ONE CUDA graph per
        # rank (keyed like the single-GPU sampler's; the counters are zeroed and the ranks meet before every replay)
        from .sampler import _graph_entry
        key = ("cfg_split", id(group), B, S, text_input_ids.shape[1], speaker_latent.shape[1], num_steps,
               tuple(plan.has_cfg), float(cfg_scale_text), float(cfg_scale_speaker), rescale_k, rescale_sigma,
               plan.uses_kv_scale)
        ent = _graph_entry(model, key)
        ent.calls += 1
        st = ent.static
        if ent.calls == 1:                                                       # first request of a shape: plain launches
            mod = model.build_mod_table(plan.t_dev)
            flags.zero_()
            torch.cuda.synchronize()
            hf.barrier()
            body(text_input_ids, text_mask, speaker_latent, speaker_mask, x_t)
        else:
            if ent.graph is None:
                st["ids"] = text_input_ids.to(dev, torch.int32).contiguous().clone()
                st["tmask"] = text_mask.to(dev).contiguous().clone()
                st["spk"] = speaker_latent.to(dev, F32).contiguous().clone()
                st["smask"] = speaker_mask.to(dev).contiguous().clone()
                st["x_t"] = x_t.clone()
                st["plan"] = plan
                mod = st["mod"] = model.build_mod_table(plan.t_dev)
                st["ws"] = model.workspace(B, S)                                   # the captured launches point into it
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g):
                    body(st["ids"], st["tmask"], st["spk"], st["smask"], st["x_t"])
                ent.graph = g
            st["ids"].copy_(text_input_ids); st["tmask"].copy_(text_mask)
            st["spk"].copy_(speaker_latent); st["smask"].copy_(speaker_mask)
            st["x_t"].copy_(x_t)
            if plan.uses_kv_scale:
                st["plan"].kv_mult.copy_(plan.kv_mult)
            st["plan"].sched.copy_(plan.sched)
            flags.zero_()
            torch.cuda.synchronize()
            hf.barrier()
            ent.graph.replay()
            x_t = st["x_t"].clone()
    if exchange == "p2p":
        torch.cuda.synchronize()
        hf.barrier()               # peers may still be reading / writing our symmetric buffers: free them together
    return x_t.view(B, S, cfg.latent_size)
