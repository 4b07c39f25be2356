"""CPU ORACLE (test infrastructure only -- never imported by the product path).

A functional fp32 restatement, in plain torch-on-CPU, of the Echo-TTS sampling hot path:
DiT forward, conditioning KV caches, the Euler/CFG samplers, PCA inverse + S1-DAC decode and
the flattening-point crop.  Every function cites the reference lines it restates
(paths are relative to the upstream repo CoreBedtime/echo-tts).  It takes a reference-keyed
``state_dict`` (see ``echo_tts_b200/schema.py``) instead of ``nn.Module`` objects so it can run
on the GPU box where the upstream sources do not exist.

Pinning: the upstream repo has no tests or golden vectors, so this file is
pinned by *executing the upstream modules themselves* in the build container:
``oracle/gen_golden.py`` imports ``/root/reference`` and writes ``tests/golden/*.pt``;
``tests/test_oracle_vs_golden.py`` checks this restatement against those fixtures, and
``tests/test_oracle_vs_reference.py`` checks it live whenever ``/root/reference`` is present.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference``
legs may import this module.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Sequence, Tuple

import torch
import torch.nn.functional as F

SD = Dict[str, torch.Tensor]
KV = List[Tuple[torch.Tensor, torch.Tensor]]


# ----------------------------------------------------------------------------------------
# small pieces
# ----------------------------------------------------------------------------------------

def rms_norm(x: torch.Tensor, w: torch.Tensor, eps: float) -> torch.Tensor:
    """model.py:99-104 (fp32 math, weight broadcast over trailing dims)."""
    xf = x.float()
    return (xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps) * w).to(x.dtype)


def rope_table(head_dim: int, n_pos: int, theta: float = 10000.0) -> Tuple[torch.Tensor, torch.Tensor]:
    """model.py:9-14 -> (cos, sin) each [n_pos, head_dim/2] fp32."""
    inv = 1.0 / (theta ** (torch.arange(0, head_dim, 2)[: head_dim // 2] / head_dim))
    ang = torch.outer(torch.arange(n_pos), inv)
    return torch.cos(ang), torch.sin(ang)


def rope_apply(x: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor) -> torch.Tensor:
    """model.py:17-24: interleaved (even, odd) pairs rotated as complex numbers.
    x: [B, S, H, Dh]; cos/sin: [S, Dh/2]."""
    xf = x.float()
    xe, xo = xf[..., 0::2], xf[..., 1::2]
    c, s = cos[None, :, None, :], sin[None, :, None, :]
    out = torch.stack([xe * c - xo * s, xe * s + xo * c], dim=-1).flatten(-2)
    return out.to(x.dtype)


def rope_first_half_heads(x: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor) -> torch.Tensor:
    """model.py:199-202: rotary on heads [0, H/2) only."""
    h = x.shape[2] // 2
    return torch.cat([rope_apply(x[:, :, :h], cos, sin), x[:, :, h:]], dim=2)


def timestep_embedding(t: torch.Tensor, size: int) -> torch.Tensor:
    """model.py:27-43: cat[cos(t w), sin(t w)], w_j = 1000 exp(-ln(1e4) j / half)."""
    half = size // 2
    w = 1000 * torch.exp(-torch.log(torch.tensor(10000.0)) * torch.arange(half, dtype=torch.float32) / half)
    a = t[..., None] * w[None]
    return torch.cat([torch.cos(a), torch.sin(a)], dim=-1).to(t.dtype)


def cond_embed(sd: SD, t: torch.Tensor, size: int) -> torch.Tensor:
    """model.py:532-538, 583: Linear-SiLU-Linear-SiLU-Linear, all bias-free -> [R, 3d]."""
    h = timestep_embedding(t, size)
    h = F.silu(F.linear(h, sd["cond_module.0.weight"]))
    h = F.silu(F.linear(h, sd["cond_module.2.weight"]))
    return F.linear(h, sd["cond_module.4.weight"])


def adaln_modulation(sd: SD, prefix: str, cond: torch.Tensor):
    """model.py:70-74, 81: (shift, scale, tanh(gate)) each [R, d]."""
    out = []
    for part, c in zip(("shift", "scale", "gate"), cond.chunk(3, dim=-1)):
        lo = F.linear(F.silu(c), sd[f"{prefix}.{part}_down.weight"])
        out.append(F.linear(lo, sd[f"{prefix}.{part}_up.weight"], sd[f"{prefix}.{part}_up.bias"]) + c)
    return out[0], out[1], torch.tanh(out[2])


def adaln_apply(x: torch.Tensor, shift: torch.Tensor, scale: torch.Tensor, eps: float) -> torch.Tensor:
    """model.py:76-79 with shift/scale [R, d] broadcast over the sequence axis."""
    xf = x.float()
    xf = xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps)
    return (xf * (scale[:, None] + 1) + shift[:, None]).to(x.dtype)


def masked_attention(q, k, v, key_mask: Optional[torch.Tensor], causal: bool = False) -> torch.Tensor:
    """softmax(q k^T / sqrt(Dh) + mask) v  (what F.scaled_dot_product_attention computes at
    model.py:148-154 / 255-261).  q: [B,S,H,Dh], k/v: [B,L,H,Dh], key_mask: [B,L] bool."""
    qh, kh, vh = (a.transpose(1, 2).float() for a in (q, k, v))
    s = qh @ kh.transpose(-1, -2) / math.sqrt(q.shape[-1])
    if key_mask is not None:
        s = s.masked_fill(~key_mask[:, None, None, :], float("-inf"))
    if causal:
        S, L = s.shape[-2:]
        s = s.masked_fill(~torch.ones(S, L, dtype=torch.bool).tril(), float("-inf"))
    return (torch.softmax(s, dim=-1) @ vh).transpose(1, 2).to(q.dtype)


def swiglu_mlp(sd: SD, prefix: str, x: torch.Tensor) -> torch.Tensor:
    """model.py:308."""
    return F.linear(F.silu(F.linear(x, sd[f"{prefix}.w1.weight"])) * F.linear(x, sd[f"{prefix}.w3.weight"]),
                    sd[f"{prefix}.w2.weight"])


# ----------------------------------------------------------------------------------------
# conditioning encoders + KV caches (once per prompt)
# ----------------------------------------------------------------------------------------

def encoder_block(sd: SD, p: str, x, heads: int, eps: float, key_mask, causal: bool, cos, sin):
    """model.py:335-339 with SelfAttention model.py:128-161 (full-head RoPE, sigmoid gate)."""
    B, S, d = x.shape
    h = rms_norm(x, sd[f"{p}.attention_norm.weight"], eps)
    q = F.linear(h, sd[f"{p}.attention.wq.weight"]).reshape(B, S, heads, -1)
    k = F.linear(h, sd[f"{p}.attention.wk.weight"]).reshape(B, S, heads, -1)
    v = F.linear(h, sd[f"{p}.attention.wv.weight"]).reshape(B, S, heads, -1)
    g = F.linear(h, sd[f"{p}.attention.gate.weight"])
    q = rope_apply(rms_norm(q, sd[f"{p}.attention.q_norm.weight"], eps), cos[:S], sin[:S])
    k = rope_apply(rms_norm(k, sd[f"{p}.attention.k_norm.weight"], eps), cos[:S], sin[:S])
    o = masked_attention(q, k, v, key_mask, causal).reshape(B, S, d) * torch.sigmoid(g)
    x = x + F.linear(o, sd[f"{p}.attention.wo.weight"])
    return x + swiglu_mlp(sd, f"{p}.mlp", rms_norm(x, sd[f"{p}.mlp_norm.weight"], eps))


def text_state(sd: SD, cfg, ids: torch.Tensor, mask: Optional[torch.Tensor]) -> torch.Tensor:
    """model.py:419-427 + text_norm (model.py:611-612)."""
    x = F.embedding(ids.long(), sd["text_encoder.text_embedding.weight"])
    cos, sin = rope_table(cfg.text_model_size // cfg.text_num_heads, ids.shape[1])
    for i in range(cfg.text_num_layers):
        x = encoder_block(sd, f"text_encoder.blocks.{i}", x, cfg.text_num_heads, cfg.norm_eps, mask, False, cos, sin)
    return rms_norm(x, sd["text_norm.weight"], cfg.norm_eps)


def patch_state(sd: SD, cfg, enc: str, norm: str, latent: torch.Tensor) -> torch.Tensor:
    """model.py:458-469 (+ speaker_norm / latent_norm): patchify(4), in_proj, /6, causal blocks."""
    B, T, C = latent.shape
    p = cfg.speaker_patch_size
    x = latent.reshape(B, T // p, C * p)
    x = F.linear(x, sd[f"{enc}.in_proj.weight"], sd[f"{enc}.in_proj.bias"]) / 6.0
    cos, sin = rope_table(cfg.speaker_model_size // cfg.speaker_num_heads, x.shape[1])
    for i in range(cfg.speaker_num_layers):
        x = encoder_block(sd, f"{enc}.blocks.{i}", x, cfg.speaker_num_heads, cfg.norm_eps, None, True, cos, sin)
    return rms_norm(x, sd[f"{norm}.weight"], cfg.norm_eps)


def _layer_kv(sd: SD, cfg, layer: int, which: str, state: torch.Tensor):
    """model.py:270-282: K = k_norm(wk_x state), V = wv_x state (decoder layer's own k_norm)."""
    B, L, _ = state.shape
    p = f"blocks.{layer}.attention"
    k = F.linear(state, sd[f"{p}.wk_{which}.weight"]).reshape(B, L, cfg.num_heads, -1)
    v = F.linear(state, sd[f"{p}.wv_{which}.weight"]).reshape(B, L, cfg.num_heads, -1)
    return rms_norm(k, sd[f"{p}.k_norm.weight"], cfg.norm_eps), v


def kv_cache_text(sd: SD, cfg, ids, mask) -> KV:
    """model.py:606-613."""
    st = text_state(sd, cfg, ids, mask)
    return [_layer_kv(sd, cfg, i, "text", st) for i in range(cfg.num_layers)]


def kv_cache_speaker(sd: SD, cfg, speaker_latent) -> KV:
    """model.py:615-621."""
    st = patch_state(sd, cfg, "speaker_encoder", "speaker_norm", speaker_latent)
    return [_layer_kv(sd, cfg, i, "speaker", st) for i in range(cfg.num_layers)]


def kv_cache_latent(sd: SD, cfg, prefix_latent) -> KV:
    """model.py:623-636, 284-293: + half-head RoPE at positions patch*4."""
    st = patch_state(sd, cfg, "latent_encoder", "latent_norm", prefix_latent)
    L = st.shape[1]
    cos, sin = rope_table(cfg.head_dim, max(L * cfg.speaker_patch_size, 1))
    pos = torch.arange(L) * cfg.speaker_patch_size
    out = []
    for i in range(cfg.num_layers):
        k, v = _layer_kv(sd, cfg, i, "latent", st)
        out.append((rope_first_half_heads(k, cos[pos], sin[pos]), v))
    return out


# ----------------------------------------------------------------------------------------
# DiT forward
# ----------------------------------------------------------------------------------------

def dit_forward(sd: SD, cfg, x, t, text_mask, speaker_mask, kv_text: KV, kv_speaker: KV,
                start_pos: Optional[int] = None, kv_latent: Optional[KV] = None) -> torch.Tensor:
    """EchoDiT.forward, model.py:563-604 (+ TransformerBlock 371-390, JointAttention 204-268)."""
    start_pos = start_pos or 0
    R, S, _ = x.shape
    H, eps = cfg.num_heads, cfg.norm_eps
    cos, sin = rope_table(cfg.head_dim, start_pos + S)
    cos, sin = cos[start_pos:], sin[start_pos:]
    spk_mask = speaker_mask[..., :: cfg.speaker_patch_size]
    cond = cond_embed(sd, t, cfg.timestep_embed_size)
    h = F.linear(x, sd["in_proj.weight"], sd["in_proj.bias"])
    for i in range(cfg.num_layers):
        p = f"blocks.{i}"
        shift, scale, gate = adaln_modulation(sd, f"{p}.attention_adaln", cond)
        hn = adaln_apply(h, shift, scale, eps)
        a = f"{p}.attention"
        q = F.linear(hn, sd[f"{a}.wq.weight"]).reshape(R, S, H, -1)
        k = F.linear(hn, sd[f"{a}.wk.weight"]).reshape(R, S, H, -1)
        v = F.linear(hn, sd[f"{a}.wv.weight"]).reshape(R, S, H, -1)
        g = F.linear(hn, sd[f"{a}.gate.weight"])
        q = rope_first_half_heads(rms_norm(q, sd[f"{a}.q_norm.weight"], eps), cos, sin)
        k = rope_first_half_heads(rms_norm(k, sd[f"{a}.k_norm.weight"], eps), cos, sin)
        ks, vs, ms = [k], [v], [torch.ones(R, S, dtype=torch.bool)]
        if kv_latent is not None and kv_latent[i][0].shape[1] > 0:
            kl, vl = kv_latent[i]
            pos = torch.arange(kl.shape[1]) * cfg.speaker_patch_size
            ks.append(kl); vs.append(vl); ms.append((pos[None] < start_pos).expand(R, -1))
        ks += [kv_text[i][0], kv_speaker[i][0]]
        vs += [kv_text[i][1], kv_speaker[i][1]]
        ms += [text_mask, spk_mask]
        o = masked_attention(q, torch.cat(ks, 1), torch.cat(vs, 1), torch.cat(ms, 1))
        o = o.reshape(R, S, -1) * torch.sigmoid(g)
        h = h + gate[:, None] * F.linear(o, sd[f"{a}.wo.weight"])
        shift, scale, gate = adaln_modulation(sd, f"{p}.mlp_adaln", cond)
        h = h + gate[:, None] * swiglu_mlp(sd, f"{p}.mlp", adaln_apply(h, shift, scale, eps))
    h = rms_norm(h, sd["out_norm.weight"], eps)
    return F.linear(h, sd["out_proj.weight"], sd["out_proj.bias"]).float()


# ----------------------------------------------------------------------------------------
# samplers
# ----------------------------------------------------------------------------------------

def t_schedule(num_steps: int) -> torch.Tensor:
    """inference.py:470,477."""
    return torch.linspace(1.0, 0.0, num_steps + 1) * 0.999


def temporal_score_rescale(v, x_t, t: float, k: float, sigma: float):
    """inference.py:431-443."""
    if t < 1:
        snr = (1 - t) ** 2 / (t ** 2)
        ratio = (snr * sigma ** 2 + 1) / (snr * sigma ** 2 / k + 1)
        return 1 / (1 - t) * (ratio * ((1 - t) * v + x_t) - x_t)
    return v


def _scale_kv(cache: KV, scale: float, max_layers: Optional[int]) -> None:
    """inference.py:420-428 (in place, first N layers, both K and V)."""
    n = len(cache) if max_layers is None else min(max_layers, len(cache))
    for i in range(n):
        cache[i][0].mul_(scale)
        cache[i][1].mul_(scale)


def _rep3(cache: KV) -> KV:
    """inference.py:409-417 specialised to three copies."""
    return [(torch.cat([k, k, k], 0), torch.cat([v, v, v], 0)) for k, v in cache]


def _euler_block(sd, cfg, x_t, sched, kv_text, kv_spk, text_mask, speaker_mask, p, trace,
                 start_pos=None, kv_lat3=None, kv_lat1=None):
    """The step loop shared by inference.py:508-558 and inference_blockwise.py:80-118."""
    B = x_t.shape[0]
    kv_text3, kv_spk3 = _rep3(kv_text), _rep3(kv_spk)
    tm3 = torch.cat([text_mask, torch.zeros_like(text_mask), text_mask], 0)
    sm3 = torch.cat([speaker_mask, speaker_mask, torch.zeros_like(speaker_mask)], 0)
    for i in range(p["num_steps"]):
        t, t_next = sched[i], sched[i + 1]
        has_cfg = bool(((t >= p["cfg_min_t"]) * (t <= p["cfg_max_t"])).item())
        if has_cfg:
            v3 = dit_forward(sd, cfg, torch.cat([x_t] * 3, 0), torch.ones(3 * B) * t, tm3, sm3,
                             kv_text3, kv_spk3, start_pos, kv_lat3)
            vc, vt, vs = v3.chunk(3, 0)
            v = vc + p["cfg_scale_text"] * (vc - vt) + p["cfg_scale_speaker"] * (vc - vs)
        else:
            v3 = dit_forward(sd, cfg, x_t, torch.ones(B) * t, text_mask, speaker_mask,
                             kv_text, kv_spk, start_pos, kv_lat1)
            v = v3
        if p["rescale_k"] is not None and p["rescale_sigma"] is not None:
            v = temporal_score_rescale(v, x_t, t, p["rescale_k"], p["rescale_sigma"])
        if p["speaker_kv_scale"] is not None and t_next < p["speaker_kv_min_t"] and t >= p["speaker_kv_min_t"]:
            _scale_kv(kv_spk, 1.0 / p["speaker_kv_scale"], p["speaker_kv_max_layers"])
            kv_spk3 = _rep3(kv_spk)
        if trace is not None:
            trace.append({"i": i, "has_cfg": has_cfg, "x_in": x_t.clone(), "v_branches": v3.clone(), "v": v.clone()})
        x_t = x_t + v * (t_next - t)
    return x_t


def sample_euler_cfg_independent_guidances(
        sd: SD, cfg, speaker_latent, speaker_mask, text_input_ids, text_mask, rng_seed: int,
        num_steps: int, cfg_scale_text: float, cfg_scale_speaker: float, cfg_min_t: float, cfg_max_t: float,
        truncation_factor=None, rescale_k=None, rescale_sigma=None, speaker_kv_scale=None,
        speaker_kv_max_layers=None, speaker_kv_min_t=None, sequence_length=None,
        noise: Optional[torch.Tensor] = None, trace: Optional[list] = None) -> torch.Tensor:
    """inference.py:446-560.  ``noise`` injects x_T (the reference draws it from a generator on the
    model's device, inference.py:475,499 -- CPU here)."""
    S = sequence_length or 640
    B = text_input_ids.shape[0]
    p = dict(num_steps=num_steps, cfg_scale_text=cfg_scale_text, cfg_scale_speaker=cfg_scale_speaker,
             cfg_min_t=cfg_min_t, cfg_max_t=cfg_max_t, rescale_k=rescale_k, rescale_sigma=rescale_sigma,
             speaker_kv_scale=speaker_kv_scale, speaker_kv_max_layers=speaker_kv_max_layers,
             speaker_kv_min_t=speaker_kv_min_t)
    kv_text = kv_cache_text(sd, cfg, text_input_ids, text_mask)
    kv_spk = kv_cache_speaker(sd, cfg, speaker_latent.float())
    if speaker_kv_scale is not None:
        _scale_kv(kv_spk, speaker_kv_scale, speaker_kv_max_layers)
    if noise is None:
        noise = torch.randn((B, S, cfg.latent_size), generator=torch.Generator().manual_seed(rng_seed))
    x_t = noise.clone().float()
    if truncation_factor is not None:
        x_t = x_t * truncation_factor
    return _euler_block(sd, cfg, x_t, t_schedule(num_steps), kv_text, kv_spk, text_mask, speaker_mask, p, trace)


def sample_blockwise_euler_cfg_independent_guidances(
        sd: SD, cfg, speaker_latent, speaker_mask, text_input_ids, text_mask, rng_seed: int,
        block_sizes: Sequence[int], num_steps: int, cfg_scale_text: float, cfg_scale_speaker: float,
        cfg_min_t: float, cfg_max_t: float, truncation_factor=None, rescale_k=None, rescale_sigma=None,
        speaker_kv_scale=None, speaker_kv_max_layers=None, speaker_kv_min_t=None,
        continuation_latent=None, noises: Optional[Sequence[torch.Tensor]] = None,
        trace: Optional[list] = None) -> torch.Tensor:
    """inference_blockwise.py:14-123."""
    B = text_input_ids.shape[0]
    p = dict(num_steps=num_steps, cfg_scale_text=cfg_scale_text, cfg_scale_speaker=cfg_scale_speaker,
             cfg_min_t=cfg_min_t, cfg_max_t=cfg_max_t, rescale_k=rescale_k, rescale_sigma=rescale_sigma,
             speaker_kv_scale=speaker_kv_scale, speaker_kv_max_layers=speaker_kv_max_layers,
             speaker_kv_min_t=speaker_kv_min_t)
    gen = torch.Generator().manual_seed(rng_seed)
    kv_text = kv_cache_text(sd, cfg, text_input_ids, text_mask)
    kv_spk = kv_cache_speaker(sd, cfg, speaker_latent.float())
    prefix = torch.zeros((B, sum(block_sizes), cfg.latent_size))
    start = 0
    if continuation_latent is not None:
        prefix = torch.cat([continuation_latent.float(), prefix], 1)
        start = continuation_latent.shape[1]
    for bi, bs in enumerate(block_sizes):
        if speaker_kv_scale is not None:                       # inference_blockwise.py:68-70 (re-applied per block)
            _scale_kv(kv_spk, speaker_kv_scale, speaker_kv_max_layers)
        kv_lat1 = kv_cache_latent(sd, cfg, prefix)             # :72-74 (whole buffer; future patches masked)
        kv_lat3 = _rep3(kv_lat1)
        x_t = noises[bi].clone().float() if noises is not None else \
            torch.randn((B, bs, cfg.latent_size), generator=gen)
        if truncation_factor is not None:
            x_t = x_t * truncation_factor
        x_t = _euler_block(sd, cfg, x_t, t_schedule(num_steps), kv_text, kv_spk, text_mask, speaker_mask, p, trace,
                           start_pos=start, kv_lat3=kv_lat3, kv_lat1=kv_lat1)
        prefix[:, start:start + bs] = x_t
        start += bs
    return prefix


# ----------------------------------------------------------------------------------------
# host-side index logic
# ----------------------------------------------------------------------------------------

def tokenizer_encode(text: str, append_bos: bool = True, normalize: bool = True) -> Tuple[List[int], str]:
    """inference.py:152-182 -> (byte ids with BOS=0, normalised text)."""
    if normalize:
        for a, b in (("…", "..."), ("’", "'"), ("”", '"'), ("\n", " "), (":", ","), (";", ","),
                     ("—", ", ")):
            text = text.replace(a, b)
        if not (text.startswith("[") or text.startswith("(") or "S1" in text or "S2" in text):
            text = "[S1] " + text
    ids = list(text.encode("utf-8"))
    if append_bos:
        ids.insert(0, 0)
    return ids, text


def text_ids_and_mask(texts: Sequence[str], max_length: Optional[int], normalize: bool = True):
    """inference.py:185-217 -> (int32 [n, L], bool [n, L], normalised texts)."""
    enc = [tokenizer_encode(t, normalize=normalize) for t in texts]
    if max_length is None:
        max_length = max(len(e) for e, _ in enc)
    ids = torch.zeros((len(texts), max_length), dtype=torch.int32)
    mask = torch.zeros((len(texts), max_length), dtype=torch.bool)
    for i, (e, _) in enumerate(enc):
        n = min(len(e), max_length)
        ids[i, :n] = torch.tensor(e[:n], dtype=torch.int32)
        mask[i, :n] = True
    return ids, mask, [t for _, t in enc]


def speaker_lengths(n_samples: int, max_latent: int = 6400, chunk: int = 640 * 2048,
                    pad_to_max: bool = False, patch: Optional[int] = 4) -> Tuple[int, int, int]:
    """Length arithmetic of inference.py:263-307 -> (n_chunks, latent_len_out, n_valid)."""
    n = min(n_samples, max_latent * 2048)
    n_chunks = -(-n // chunk) if n > 0 else 0
    total = n_chunks * (chunk // 2048)
    actual = n // 2048
    if pad_to_max and total < max_latent:
        total = max_latent
    elif not pad_to_max:
        total = actual
    if patch is not None:
        total = total // patch * patch
    return n_chunks, total, min(actual, total)


def find_flattening_point(data: torch.Tensor, target_value=0.0, window_size=20, std_threshold=0.05) -> int:
    """inference.py:315-330 (data [length, 80]; unbiased std over the whole window)."""
    pad = torch.cat([data, torch.zeros(window_size, *data.shape[1:], dtype=data.dtype)])
    for i in range(len(pad) - window_size):
        w = pad[i:i + window_size]
        if w.std() < std_threshold and abs(w.mean() - target_value) < 0.1:
            return i
    return len(data)


# ----------------------------------------------------------------------------------------
# PCA inverse + Fish S1-DAC decode (decode_zq)
# ----------------------------------------------------------------------------------------

def wn_weight(sd: SD, prefix: str) -> torch.Tensor:
    """weight_norm parametrization (autoencoder.py:90-94, 291-293): w = g * v / ||v||, the norm
    taken over every dim but 0."""
    if f"{prefix}.conv.parametrizations.weight.original0" in sd:
        prefix = f"{prefix}.conv"                         # CausalConvNet wrapper; VectorQuantize's WNConv1d has none
    g = sd[f"{prefix}.parametrizations.weight.original0"]
    v = sd[f"{prefix}.parametrizations.weight.original1"]
    return g * v / v.flatten(1).norm(dim=1).reshape(-1, *([1] * (v.dim() - 1)))


def causal_conv1d(x, w, b, dilation: int = 1, groups: int = 1):
    """CausalConvNet.forward, autoencoder.py:285-289, stride 1: left pad (k-1)*d zeros."""
    k = w.shape[-1]
    return F.conv1d(F.pad(x, ((k - 1) * dilation, 0)), w, b, dilation=dilation, groups=groups)


def causal_conv_transpose1d(x, w, b, stride: int):
    """CausalTransConvNet.forward, autoencoder.py:310-316: trim (k - stride) samples on the right."""
    y = F.conv_transpose1d(x, w, b, stride=stride)
    trim = w.shape[-1] - stride
    return y[..., : y.shape[-1] - trim] if trim > 0 else y


def snake(x, alpha):
    """autoencoder.py:96-102."""
    return x + (alpha + 1e-9).reciprocal() * torch.sin(alpha * x).pow(2)


def ae_rms_norm(x, w, eps):
    """autoencoder.py:726-731 (normalise in fp32, cast back, *then* multiply by weight)."""
    xf = x.float()
    return (xf * torch.rsqrt((xf * xf).mean(-1, keepdim=True) + eps)).type_as(x) * w


def ae_window_transformer(sd: SD, cfg, x: torch.Tensor, pm: str = "quantizer.post_module", layers: Optional[int] = None,
                          heads: Optional[int] = None, window: Optional[int] = None) -> torch.Tensor:
    """WindowLimitedTransformer.forward, autoencoder.py:786-802 (channels-first in/out) with
    Transformer/TransformerBlock/Attention/FeedForward 590-717 and the window mask 762-773.
    RoPE table is the module's bf16 buffer (autoencoder.py:805-826), applied to ALL heads."""
    h = x.transpose(1, 2)                                     # [B, T, D]
    B, T, D = h.shape
    H, Dh = heads or cfg.tr_heads, cfg.tr_head_dim
    window = window or cfg.tr_window
    fc = sd[f"{pm}.freqs_cis"][:T].float()                    # [T, Dh/2, 2]
    cos, sin = fc[..., 0], fc[..., 1]
    i = torch.arange(T)
    win = (i[None, :] <= i[:, None]) & (i[None, :] >= (i[:, None] - window + 1).clamp(min=0))
    for l in range(cfg.tr_layers if layers is None else layers):
        p = f"{pm}.layers.{l}"
        a = ae_rms_norm(h, sd[f"{p}.attention_norm.weight"], cfg.tr_eps)
        q, k, v = F.linear(a, sd[f"{p}.attention.wqkv.weight"]).split(H * Dh, dim=-1)
        q = rope_apply(q.reshape(B, T, H, Dh), cos, sin).transpose(1, 2)
        k = rope_apply(k.reshape(B, T, H, Dh), cos, sin).transpose(1, 2)
        v = v.reshape(B, T, H, Dh).transpose(1, 2)
        s = (q @ k.transpose(-1, -2)) / math.sqrt(Dh)
        s = s.masked_fill(~win[None, None], float("-inf"))
        o = (torch.softmax(s, -1) @ v).transpose(1, 2).reshape(B, T, H * Dh)
        h = h + F.linear(o, sd[f"{p}.attention.wo.weight"]) * sd[f"{p}.attention_layer_scale.gamma"]
        f = ae_rms_norm(h, sd[f"{p}.ffn_norm.weight"], cfg.tr_eps)
        f = F.linear(F.silu(F.linear(f, sd[f"{p}.feed_forward.w1.weight"])) *
                     F.linear(f, sd[f"{p}.feed_forward.w3.weight"]), sd[f"{p}.feed_forward.w2.weight"])
        h = h + f * sd[f"{p}.ffn_layer_scale.gamma"]
    h = ae_rms_norm(h, sd[f"{pm}.norm.weight"], cfg.tr_eps)
    return h.transpose(1, 2)


def convnext_block(sd: SD, p: str, x: torch.Tensor) -> torch.Tensor:
    """ConvNeXtBlock.forward, autoencoder.py:360-373: dwconv k7 -> LayerNorm(1e-6) -> Linear -> GELU(erf) -> Linear ->
    gamma -> + residual."""
    y = causal_conv1d(x, sd[f"{p}.dwconv.conv.weight"], sd[f"{p}.dwconv.conv.bias"], groups=x.shape[1])
    y = y.transpose(1, 2)
    y = F.layer_norm(y, (y.shape[-1],), sd[f"{p}.norm.weight"], sd[f"{p}.norm.bias"], 1e-6)
    y = F.linear(F.gelu(F.linear(y, sd[f"{p}.pwconv1.weight"], sd[f"{p}.pwconv1.bias"])),
                 sd[f"{p}.pwconv2.weight"], sd[f"{p}.pwconv2.bias"])
    return x + (y * sd[f"{p}.gamma"]).transpose(1, 2)


def ae_upsample(sd: SD, cfg, x: torch.Tensor) -> torch.Tensor:
    """quantizer.upsample, autoencoder.py:427-435: per factor, causal transposed conv then ConvNeXt."""
    for u, f in enumerate(cfg.up_factors):
        p = f"quantizer.upsample.{u}"
        x = causal_conv_transpose1d(x, sd[f"{p}.0.conv.weight"], sd[f"{p}.0.conv.bias"], f)
        x = convnext_block(sd, f"{p}.1", x)
    return x


def ae_wave_decoder(sd: SD, cfg, x: torch.Tensor) -> torch.Tensor:
    """Decoder.forward, autoencoder.py:971-998 with DecoderBlock 959-965 and ResidualUnit 884-900
    (causal; the DecoderBlock transformer is constructed but never run, 944-965)."""
    dm = "decoder.model"
    x = causal_conv1d(x, wn_weight(sd, f"{dm}.0"), sd[f"{dm}.0.conv.bias"])
    for i, r in enumerate(cfg.decoder_rates):
        b = f"{dm}.{i + 1}.block"
        x = snake(x, sd[f"{b}.0.alpha"])
        x = causal_conv_transpose1d(x, wn_weight(sd, f"{b}.1"), sd[f"{b}.1.conv.bias"], r)
        for j, dil in enumerate((1, 3, 9)):
            ru = f"{b}.{2 + j}.block"
            y = snake(x, sd[f"{ru}.0.alpha"])
            y = causal_conv1d(y, wn_weight(sd, f"{ru}.1"), sd[f"{ru}.1.conv.bias"], dilation=dil)
            y = snake(y, sd[f"{ru}.2.alpha"])
            y = causal_conv1d(y, wn_weight(sd, f"{ru}.3"), sd[f"{ru}.3.conv.bias"])
            x = x + y
    n = len(cfg.decoder_rates)
    x = snake(x, sd[f"{dm}.{n + 1}.alpha"])
    x = causal_conv1d(x, wn_weight(sd, f"{dm}.{n + 2}"), sd[f"{dm}.{n + 2}.conv.bias"])
    return torch.tanh(x)


def decode_zq(sd: SD, cfg, z: torch.Tensor) -> torch.Tensor:
    """DAC.decode_zq, autoencoder.py:1128-1132.  z: [B, latent_dim, T] -> [B, 1, T*hop]."""
    return ae_wave_decoder(sd, cfg, ae_upsample(sd, cfg, ae_window_transformer(sd, cfg, z)))


def ae_decode(sd: SD, cfg, pca_components, pca_mean, latent_scale: float, z: torch.Tensor) -> torch.Tensor:
    """inference.py:232-235.  z: [B, T, 80] fp32 -> waveform [B, 1, T*hop] fp32."""
    z = (z / latent_scale) @ pca_components + pca_mean
    return decode_zq(sd, cfg, z.transpose(1, 2)).float()


def synth_pca(latent_size: int = 80, latent_dim: int = 1024, seed: int = 99):
    """Synthetic PCAState constants: lives in ``echo_tts_b200.schema`` (shared with the product-side bench arm)."""
    from echo_tts_b200.schema import synth_pca as _sp
    return _sp(latent_size, latent_dim, seed)


# ---------------------------------------------------------------------------------------------------------------------
# LoRA adapters: the un-merged forward of a wrapped layer, lora.py:55-64, and the merge lora.py:66-82
# ---------------------------------------------------------------------------------------------------------------------
def lora_linear(x: torch.Tensor, w: torch.Tensor, a: torch.Tensor, b: torch.Tensor, alpha: float, rank: int) -> torch.Tensor:
    """y = x W^T + (alpha / rank) * (x A^T) B^T   (LoRALinear.forward, lora.py:55-64; dropout is identity at inference)"""
    return x @ w.T + ((x @ a.T) @ b.T) * (alpha / rank)


def lora_merged_weight(w: torch.Tensor, a: torch.Tensor, b: torch.Tensor, alpha: float, rank: int) -> torch.Tensor:
    """W' = W + (alpha / rank) * B A   (LoRALinear.merge_weights, lora.py:77-78)"""
    return w + (b @ a) * (alpha / rank)


# ---------------------------------------------------------------------------------------------------------------------
# Fish S1-DAC encode: DAC.encode_zq, autoencoder.py:1080-1126
# ---------------------------------------------------------------------------------------------------------------------
def causal_conv1d_strided(x, w, b, stride: int):
    """CausalConvNet.forward with stride, autoencoder.py:285-289 + get_extra_padding_for_conv1d :49-56: left pad
    k - stride zeros, right pad up to a whole number of frames."""
    k, length = w.shape[-1], x.shape[-1]
    pad = k - stride
    n_frames = (length - k + pad) / stride + 1
    extra = (math.ceil(n_frames) - 1) * stride + (k - pad) - length
    return F.conv1d(F.pad(x, (pad, extra)), w, b, stride=stride)


def residual_unit(sd: SD, ru: str, x: torch.Tensor, dilation: int) -> torch.Tensor:
    """ResidualUnit.forward (causal), autoencoder.py:879-900."""
    y = snake(x, sd[f"{ru}.0.alpha"])
    y = causal_conv1d(y, wn_weight(sd, f"{ru}.1"), sd[f"{ru}.1.conv.bias"], dilation=dilation)
    y = snake(y, sd[f"{ru}.2.alpha"])
    y = causal_conv1d(y, wn_weight(sd, f"{ru}.3"), sd[f"{ru}.3.conv.bias"])
    return x + y


def ae_encoder(sd: SD, cfg, audio: torch.Tensor) -> torch.Tensor:
    """Encoder.forward, autoencoder.py:903-929 with EncoderBlock 839-876: audio [B, 1, n] -> z [B, latent_dim, n / prod(rates)]."""
    eb = "encoder.block"
    x = causal_conv1d(audio, wn_weight(sd, f"{eb}.0"), sd[f"{eb}.0.conv.bias"])
    n = len(cfg.encoder_rates)
    for i, r in enumerate(cfg.encoder_rates):
        b = f"{eb}.{i + 1}.block"
        for j, dil in enumerate((1, 3, 9)):
            x = residual_unit(sd, f"{b}.{j}.block", x, dil)
        x = snake(x, sd[f"{b}.3.alpha"])
        x = causal_conv1d_strided(x, wn_weight(sd, f"{b}.4"), sd[f"{b}.4.conv.bias"], r)
        if i == n - 1 and cfg.enc_tr_layers > 0:
            x = ae_window_transformer(sd, cfg, x, f"{b}.5", layers=cfg.enc_tr_layers, heads=cfg.enc_tr_heads,
                                      window=cfg.enc_tr_window)
    x = snake(x, sd[f"{eb}.{n + 1}.alpha"])
    return causal_conv1d(x, wn_weight(sd, f"{eb}.{n + 2}"), sd[f"{eb}.{n + 2}.conv.bias"])


def ae_downsample(sd: SD, cfg, z: torch.Tensor) -> torch.Tensor:
    """quantizer.downsample, autoencoder.py:416-425: per factor, strided causal conv (k = stride) then ConvNeXt."""
    for u, f in enumerate(cfg.up_factors):
        p = f"quantizer.downsample.{u}"
        z = causal_conv1d_strided(z, sd[f"{p}.0.conv.weight"], sd[f"{p}.0.conv.bias"], f)
        z = convnext_block(sd, f"{p}.1", z)
    return z


def vq_nearest(e: torch.Tensor, codebook: torch.Tensor) -> torch.Tensor:
    """VectorQuantize.decode_latents, autoencoder.py:145-157: e [B, dim, T] -> indices [B, T] of the nearest
    l2-normalised code (first index on exact ties, as torch.max)."""
    B = e.shape[0]
    enc = F.normalize(e.transpose(1, 2).reshape(-1, e.shape[1]))
    cb = F.normalize(codebook)
    dist = enc.pow(2).sum(1, keepdim=True) - 2 * enc @ cb.t() + cb.pow(2).sum(1, keepdim=True).t()
    return (-dist).max(1)[1].reshape(B, -1)


def vq_quantize(sd: SD, p: str, z: torch.Tensor):
    """VectorQuantize.forward (eval), autoencoder.py:129-136: -> (out_proj(code vectors) [B, D, T], indices [B, T])."""
    z_e = F.conv1d(z, wn_weight(sd, f"{p}.in_proj"), sd[f"{p}.in_proj.bias"])
    idx = vq_nearest(z_e, sd[f"{p}.codebook.weight"])
    z_q = F.embedding(idx, sd[f"{p}.codebook.weight"]).transpose(1, 2)
    z_q = z_e + (z_q - z_e)                                  # straight-through form the reference evaluates
    return F.conv1d(z_q, wn_weight(sd, f"{p}.out_proj"), sd[f"{p}.out_proj.bias"]), idx


def rvq_codes(sd: SD, cfg, z: torch.Tensor) -> torch.Tensor:
    """Semantic VQ then the residual stack (DownsampleResidualVectorQuantize.forward :457-468 with
    ResidualVectorQuantize.forward :185-222, eval mode): z [B, D, T] -> codes [B, 1 + n_codebooks, T] int64."""
    sem, idx0 = vq_quantize(sd, "quantizer.semantic_quantizer.quantizers.0", z)
    residual = z - sem
    codes = [idx0]
    for q in range(cfg.n_codebooks):
        z_q, idx = vq_quantize(sd, f"quantizer.quantizer.quantizers.{q}", residual)
        residual = residual - z_q
        codes.append(idx)
    return torch.stack(codes, 1)


def zq_from_codes(sd: SD, cfg, codes: torch.Tensor) -> torch.Tensor:
    """encode_zq's second half, autoencoder.py:1116-1126 with from_codes :224-233: clamp, look up, out_proj, sum."""
    z_q = 0.0
    for i in range(codes.shape[1]):
        p = "quantizer.semantic_quantizer.quantizers.0" if i == 0 else f"quantizer.quantizer.quantizers.{i - 1}"
        cb = sd[f"{p}.codebook.weight"]
        v = F.embedding(codes[:, i].clamp(max=cb.shape[0] - 1), cb).transpose(1, 2)
        z_q = z_q + F.conv1d(v, wn_weight(sd, f"{p}.out_proj"), sd[f"{p}.out_proj.bias"])
    return z_q


def encode_stages(sd: SD, cfg, audio: torch.Tensor) -> dict:
    """DAC.encode (autoencoder.py:1080-1100) + encode_zq (:1116-1126) with the intermediate tensors the stage-wise
    parity tests pin: audio [B, 1, n] or [B, n] -> {z_enc, z_pre (quantizer input after pre_module), codes, z_q}."""
    if audio.dim() == 2:
        audio = audio.unsqueeze(1)
    n = audio.shape[-1]
    audio = F.pad(audio, (0, math.ceil(n / cfg.hop) * cfg.hop - n))
    z_enc = ae_encoder(sd, cfg, audio)
    z_pre = ae_window_transformer(sd, cfg, ae_downsample(sd, cfg, z_enc), "quantizer.pre_module")
    codes = rvq_codes(sd, cfg, z_pre)
    return dict(z_enc=z_enc, z_pre=z_pre, codes=codes, z_q=zq_from_codes(sd, cfg, codes))


def encode_zq(sd: SD, cfg, audio: torch.Tensor) -> torch.Tensor:
    """DAC.encode_zq, autoencoder.py:1116-1126: audio -> z_q [B, latent_dim, ceil(n / hop)]."""
    return encode_stages(sd, cfg, audio)["z_q"]


def ae_encode(sd: SD, cfg, pca_components, pca_mean, latent_scale: float, audio: torch.Tensor) -> torch.Tensor:
    """inference.py:223-229: z_q -> PCA projection: ((z_q^T - mean) @ components^T) * latent_scale, [B, T, 80]."""
    z_q = encode_zq(sd, cfg, audio).float()
    return ((z_q.transpose(1, 2) - pca_mean) @ pca_components.T) * latent_scale


# ---------------------------------------------------------------------------------------------------------------------
# Audio front-end: load_audio after the decode, inference.py:145-147.  The resampler is third party:
# torchaudio (requirements.txt; 2.11.0 installed here) functional.resample, "sinc_interp_hann", lowpass_filter_width 6,
# rolloff 0.99 -- restated from its published algorithm (_get_sinc_resample_kernel / _apply_sinc_resample_kernel) and
# pinned against the installed torchaudio in tests/test_audio_frontend.py.
# ---------------------------------------------------------------------------------------------------------------------
def resample_sinc(waveform: torch.Tensor, orig_freq: int, new_freq: int, lowpass_filter_width: int = 6,
                  rolloff: float = 0.99) -> torch.Tensor:
    """waveform fp32 [..., n] -> [..., ceil(n * new / orig)]: y[f*new + p] = sum_k bank[p, k] * xpad[f*orig + k] with the
    Hann-windowed sinc bank of width ceil(lpw * orig / (min(orig, new) * rolloff)); frequencies reduced by their gcd."""
    if orig_freq == new_freq:
        return waveform
    g = math.gcd(int(orig_freq), int(new_freq))
    orig, new = int(orig_freq) // g, int(new_freq) // g
    base = min(orig, new) * rolloff
    width = math.ceil(lowpass_filter_width * orig / base)
    dt = waveform.dtype
    idx = torch.arange(-width, width + orig, dtype=dt)[None, None] / orig
    t = torch.arange(0, -new, -1, dtype=dt)[:, None, None] / new + idx
    t = (t * base).clamp(-lowpass_filter_width, lowpass_filter_width)
    window = torch.cos(t * math.pi / lowpass_filter_width / 2) ** 2
    t = t * math.pi
    bank = torch.where(t == 0, torch.ones((), dtype=dt), t.sin() / t) * (window * (base / orig))    # [new, 1, taps]
    shape = waveform.shape
    x = waveform.reshape(-1, shape[-1])
    n = x.shape[1]
    x = F.pad(x, (width, width + orig))
    y = F.conv1d(x[:, None], bank, stride=orig).transpose(1, 2).reshape(x.shape[0], -1)
    y = y[:, : math.ceil(new * n / orig)]
    return y.reshape(shape[:-1] + y.shape[-1:])


def load_audio_post(samples: torch.Tensor, sample_rate: int) -> torch.Tensor:
    """inference.py:145-147 on decoded samples [channels, n]: channel mean, resample to 44.1 kHz, / max(peak, 1)."""
    audio = samples.mean(dim=0).unsqueeze(0)
    audio = resample_sinc(audio, sample_rate, 44_100)
    return audio / torch.maximum(audio.abs().max(), torch.tensor(1.0))


# ---------------------------------------------------------------------------------------------------------------------
# sample_pipeline (inference.py:346-400) and get_speaker_latent_and_mask (inference.py:250-309) composed from the
# restatements above.  Pinned live against the upstream functions in tests/test_oracle_vs_reference.py.
# ---------------------------------------------------------------------------------------------------------------------
def speaker_latent_and_mask(asd: SD, acfg, pca, audio: torch.Tensor, max_speaker_latent_length: int = 6400,
                            audio_chunk_size: int = 640 * 2048, pad_to_max: bool = False,
                            divis_by_patch_size: Optional[int] = 4):
    """inference.py:250-309: audio [1, n] -> (latent fp32 [1, T, 80], mask bool [1, T]); the S1-DAC hop of 2048 is
    hard-coded upstream (:262), so this holds for the full-size autoencoder only."""
    comp, mean, scale = pca
    hop = 2048
    assert audio.ndim == 2 and audio.shape[0] == 1
    audio = audio[:, : max_speaker_latent_length * hop]
    chunks = []
    for i in range(0, audio.shape[1], audio_chunk_size):
        c = audio[:, i:i + audio_chunk_size]
        if c.shape[1] < audio_chunk_size:
            c = F.pad(c, (0, audio_chunk_size - c.shape[1]))
        chunks.append(ae_encode(asd, acfg, comp, mean, scale, c.unsqueeze(0)))
    lat = torch.cat(chunks, dim=1)
    actual = audio.shape[1] // hop
    mask = (torch.arange(lat.shape[1]) < actual).unsqueeze(0)
    if pad_to_max and lat.shape[1] < max_speaker_latent_length:
        lat = F.pad(lat, (0, 0, 0, max_speaker_latent_length - lat.shape[1]))
        mask = F.pad(mask, (0, max_speaker_latent_length - mask.shape[1]))
    elif not pad_to_max:
        lat, mask = lat[:, :actual], mask[:, :actual]
    if divis_by_patch_size is not None:
        n = lat.shape[1] // divis_by_patch_size * divis_by_patch_size
        lat, mask = lat[:, :n], mask[:, :n]
    return lat, mask


def sample_pipeline(sd: SD, cfg, asd: SD, acfg, pca, sampler_kwargs: dict, text_prompt: str, speaker_audio, rng_seed: int,
                    pad_to_max_speaker_latent_length: Optional[int] = None, pad_to_max_text_length: Optional[int] = None,
                    normalize_text: bool = False, noise: Optional[torch.Tensor] = None, hop: int = 2048):
    """inference.py:346-400 -> (audio fp32 [1, 1, crop * hop], normalised text, latent).  ``speaker_audio``: None (the
    zero speaker of :377-383), a waveform [1, n], or an already encoded (latent, mask) pair.  The text is ALWAYS padded to
    max_length = min(pad_to_max_text_length or 768, 768): get_text_input_ids_and_mask allocates max_length columns and its
    ``pad_to_max=False`` slice ``[:, :max_length]`` is a no-op (:209-210).  ``hop`` = samples per latent (2048 upstream,
    hard-coded in the crop :337)."""
    max_len = min(pad_to_max_text_length or 768, 768)
    ids, mask, texts = text_ids_and_mask([text_prompt], max_len, normalize=normalize_text)
    if speaker_audio is None:
        n = pad_to_max_speaker_latent_length or 4
        spk, smask = torch.zeros((1, n, 80)), torch.zeros((1, n), dtype=torch.bool)
    elif isinstance(speaker_audio, tuple):
        spk, smask = speaker_audio
    else:
        spk, smask = speaker_latent_and_mask(asd, acfg, pca, speaker_audio,
                                             max_speaker_latent_length=pad_to_max_speaker_latent_length or 6400,
                                             pad_to_max=(pad_to_max_speaker_latent_length is not None))
    latent = sample_euler_cfg_independent_guidances(sd, cfg, spk, smask, ids, mask, rng_seed, noise=noise, **sampler_kwargs)
    comp, mean, scale = pca
    audio = ae_decode(asd, acfg, comp, mean, scale, latent)
    crop = find_flattening_point(latent[0])
    return audio[..., : crop * hop], texts[0], latent
