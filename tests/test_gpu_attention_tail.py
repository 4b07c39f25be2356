"""Attention, sampler tail and small kernels on a B200 against torch fp32 references."""
import math

import pytest
import torch

from conftest import rel_l2

pytestmark = pytest.mark.gpu
BF16, F32 = torch.bfloat16, torch.float32


@pytest.fixture(scope="module")
def ops():
    from echo_tts_b200 import ops as o
    return o


def _rand(shape, seed, scale=1.0, dtype=F32):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return (torch.randn(shape, device="cuda", generator=g) * scale).to(dtype)


def _ref_attention(q, segs, gate):
    """fp32 dense reference: concat segments (what the reference model does, model.py:246-261)."""
    R, S, H, Dh = q.shape
    ks, vs, ms = [], [], []
    for s in segs:
        Lk = s["k"].shape[1]
        idx = torch.arange(R, device=q.device) % s["k"].shape[0]
        f = s.get("scale", 1.0)
        ks.append(s["k"].float()[idx] * f)
        vs.append(s["v"].float()[idx] * f)
        m = s["mask"].bool() if s.get("mask") is not None else torch.ones(R, Lk, dtype=torch.bool, device=q.device)
        m = m[:, None, :].expand(R, S, Lk).clone()
        if s.get("causal"):
            i = torch.arange(S, device=q.device)[:, None]
            j = torch.arange(Lk, device=q.device)[None, :]
            c = j <= i
            if s.get("window", 0) > 0:
                c &= j > i - s["window"]
            m &= c[None]
        ms.append(m)
    K, V, Mk = torch.cat(ks, 1), torch.cat(vs, 1), torch.cat(ms, 2)
    sc = torch.einsum("rshd,rlhd->rhsl", q.float(), K) / math.sqrt(Dh)
    sc = sc.masked_fill(~Mk[:, None], float("-inf"))
    o = torch.einsum("rhsl,rlhd->rshd", torch.softmax(sc, -1), V).reshape(R * S, H * Dh)
    return o * gate.float() if gate is not None else o


def test_joint_attention_segments(ops):
    """CFG layout: 3B rows share B-row caches; dropped modalities via zero masks; speaker K/V scaled."""
    B, S, H, Dh, Lt, Ls, Lp = 2, 200, 3, 128, 150, 37, 20
    R = 3 * B
    q, k, v = (_rand((R, S, H, Dh), i, dtype=BF16) for i in (1, 2, 3))
    kt, vt, ksp, vsp, kl, vl = (_rand((B, n, H, Dh), 4 + i, dtype=BF16) for i, n in enumerate((Lt, Lt, Ls, Ls, Lp, Lp)))
    gate = torch.sigmoid(_rand((R * S, H * Dh), 10)).to(BF16)
    tm = torch.zeros(B, Lt, dtype=torch.uint8, device="cuda"); tm[0, :70] = 1; tm[1, :133] = 1; tm[1, 5] = 0
    sm = torch.ones(B, Ls, dtype=torch.uint8, device="cuda"); sm[1, 20:] = 0
    tm3 = torch.cat([tm, torch.zeros_like(tm), tm]); sm3 = torch.cat([sm, sm, torch.zeros_like(sm)])
    scale = torch.tensor([1.5], device="cuda")
    n_prefix = 13
    out = torch.empty((R * S, H * Dh), device="cuda", dtype=BF16)
    ops.attention(q, [ops.Segment(k, v), ops.Segment(kl, vl, kv_batch_mod=B, length=n_prefix),
                      ops.Segment(kt, vt, kv_batch_mod=B, mask=tm3, extent=ops.mask_extent(tm3)),
                      ops.Segment(ksp, vsp, kv_batch_mod=B, mask=sm3, extent=ops.mask_extent(sm3), kv_scale=scale)],
                  out, gate=gate)
    ref = _ref_attention(q, [dict(k=k, v=v), dict(k=kl[:, :n_prefix], v=vl[:, :n_prefix]), dict(k=kt, v=vt, mask=tm3),
                             dict(k=ksp, v=vsp, mask=sm3, scale=1.5)], gate)
    assert rel_l2(out.float(), ref) < 6e-3


@pytest.mark.parametrize("impl", [0, 1])
@pytest.mark.parametrize("R,S,Lt,valid_t,Ls", [(3, 640, 768, 256, 53), (1, 640, 768, 300, 53), (2, 128, 64, 64, 8),
                                                (3, 320, 768, 1, 1600)])
def test_joint_attention_dit_shapes(ops, impl, R, S, Lt, valid_t, Ls):
    """BASELINE config-2/4 shapes: self + text (768-padded) + speaker segments, CFG rows dropping a modality,
    tcgen05 kernel (impl 0) and mma.sync kernel (impl 1) against the dense fp32 reference."""
    H, Dh = 16, 128
    qkv = _rand((3, R, S, H, Dh), 31, dtype=BF16)
    kt, vt = _rand((1, Lt, H, Dh), 32, dtype=BF16), _rand((1, Lt, H, Dh), 33, dtype=BF16)
    ksp, vsp = _rand((1, Ls, H, Dh), 34, dtype=BF16), _rand((1, Ls, H, Dh), 35, dtype=BF16)
    gate = torch.sigmoid(_rand((R * S, H * Dh), 36)).to(BF16)
    tm = torch.zeros(R, Lt, dtype=torch.uint8, device="cuda"); tm[:, :valid_t] = 1
    sm = torch.ones(R, Ls, dtype=torch.uint8, device="cuda")
    if R == 3:
        tm[1] = 0; sm[2] = 0
    scale = torch.tensor([1.5], device="cuda")
    out = torch.full((R * S, H * Dh), float("nan"), device="cuda", dtype=BF16)
    ops.attention(qkv[0], [ops.Segment(qkv[1], qkv[2]),
                           ops.Segment(kt, vt, kv_batch_mod=1, mask=tm, extent=ops.mask_extent(tm)),
                           ops.Segment(ksp, vsp, kv_batch_mod=1, mask=sm, extent=ops.mask_extent(sm), kv_scale=scale)],
                  out, gate=gate, impl=impl)
    ref = _ref_attention(qkv[0], [dict(k=qkv[1], v=qkv[2]), dict(k=kt, v=vt, mask=tm), dict(k=ksp, v=vsp, mask=sm, scale=1.5)],
                         gate)
    assert rel_l2(out.float(), ref) < 6e-3


@pytest.mark.parametrize("splits", [0, 2, 3, 4])
@pytest.mark.parametrize("R,S,Lt,valid_t,Ls,Lp", [(1, 640, 768, 256, 53, 0),      # un-guided 640-token step: 80 groups
                                                   (3, 160, 768, 256, 53, 480),    # guided 160-latent block behind 480
                                                   (1, 160, 768, 0, 0, 0),         # self keys only: 3 tiles, 4 parts asked
                                                   (1, 100, 64, 1, 5, 0)])         # two tiles in all
def test_split_kv_attention(ops, splits, R, S, Lt, valid_t, Ls, Lp):
    """Split-KV (several CTAs share one query tile's keys, partial (O, m, l) merged in part order by the last to arrive):
    equal to the dense fp32 reference at the tolerance of the unsplit kernel, close to the unsplit kernel, bit-identical
    from run to run (the merge order is fixed), counters self-resetting (three launches in a row)."""
    H, Dh = 16, 128
    qkv = _rand((3, R, S, H, Dh), 41, dtype=BF16)
    kt, vt = _rand((1, Lt, H, Dh), 42, dtype=BF16), _rand((1, Lt, H, Dh), 43, dtype=BF16)
    ksp, vsp = _rand((1, max(Ls, 1), H, Dh), 44, dtype=BF16), _rand((1, max(Ls, 1), H, Dh), 45, dtype=BF16)
    kp, vp = _rand((1, max(Lp, 1), H, Dh), 46, dtype=BF16), _rand((1, max(Lp, 1), H, Dh), 47, dtype=BF16)
    gate = torch.sigmoid(_rand((R * S, H * Dh), 48)).to(BF16)
    tm = torch.zeros(R, Lt, dtype=torch.uint8, device="cuda"); tm[:, :valid_t] = 1
    sm = torch.ones(R, max(Ls, 1), dtype=torch.uint8, device="cuda")
    if Ls == 0:
        sm[:] = 0
    if R == 3:
        tm[1] = 0; sm[2] = 0
    segs = [ops.Segment(qkv[1], qkv[2])]
    refs = [dict(k=qkv[1], v=qkv[2])]
    if Lp:
        segs.append(ops.Segment(kp, vp, kv_batch_mod=1, length=Lp)); refs.append(dict(k=kp, v=vp))
    segs += [ops.Segment(kt, vt, kv_batch_mod=1, mask=tm, extent=ops.mask_extent(tm)),
             ops.Segment(ksp, vsp, kv_batch_mod=1, mask=sm, extent=ops.mask_extent(sm))]
    refs += [dict(k=kt, v=vt, mask=tm), dict(k=ksp, v=vsp, mask=sm)]
    outs = []
    for _ in range(3):
        out = torch.full((R * S, H * Dh), float("nan"), device="cuda", dtype=BF16)
        ops.attention(qkv[0], segs, out, gate=gate, kv_splits=splits)
        outs.append(out)
    base = torch.empty_like(outs[0])
    ops.attention(qkv[0], segs, base, gate=gate, kv_splits=1)
    ref = _ref_attention(qkv[0], refs, gate)
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[1], outs[2])
    assert rel_l2(outs[0].float(), ref) < 6e-3 and rel_l2(base.float(), ref) < 6e-3
    assert rel_l2(outs[0].float(), base.float()) < 4e-3


@pytest.mark.parametrize("Dh,window", [(128, 0), (64, 16), (64, 128)])
def test_causal_window_attention(ops, Dh, window):
    B, S, H = 2, 333, 4
    q, k, v = (_rand((B, S, H, Dh), i, dtype=BF16) for i in (11, 12, 13))
    out = torch.empty((B * S, H * Dh), device="cuda", dtype=BF16)
    ops.attention(q, [ops.Segment(k, v, causal=True, window=window)], out)
    ref = _ref_attention(q, [dict(k=k, v=v, causal=True, window=window)], None)
    assert rel_l2(out.float(), ref) < 6e-3


@pytest.mark.parametrize("S,off", [(333, 0), (128, 0), (1600, 0), (100, 480), (64, 1), (200, 63)])
@pytest.mark.parametrize("impl", [0, 1])
def test_causal_encoder_attention_with_query_offset(ops, impl, S, off):
    """The causal speaker / latent-prefix encoders' attention (head_dim 128; 1600 patches = a 6400-latent reference; queries
    that are the LAST S positions of off + S keys = the incremental prefix encoder): tcgen05 kernel (impl 0: tiles past the
    diagonal skipped, tiles on it masked per row) and mma.sync kernel (impl 1) against the dense fp32 reference, gated."""
    B, H, Dh = 2, 4, 128
    q = _rand((B, S, H, Dh), 61, dtype=BF16)
    k, v = _rand((B, off + S, H, Dh), 62, dtype=BF16), _rand((B, off + S, H, Dh), 63, dtype=BF16)
    gate = torch.sigmoid(_rand((B * S, H * Dh), 64)).to(BF16)
    out = torch.full((B * S, H * Dh), float("nan"), device="cuda", dtype=BF16)
    ops.attention(q, [ops.Segment(k, v, causal=True, q_offset=off)], out, gate=gate, impl=impl)
    sc = torch.einsum("rshd,rlhd->rhsl", q.float(), k.float()) / math.sqrt(Dh)
    i, j = torch.arange(S, device="cuda")[:, None], torch.arange(off + S, device="cuda")[None, :]
    sc = sc.masked_fill(~(j <= i + off)[None, None], float("-inf"))
    ref = torch.einsum("rhsl,rlhd->rshd", torch.softmax(sc, -1), v.float()).reshape(B * S, H * Dh) * gate.float()
    assert rel_l2(out.float(), ref) < 6e-3


def test_key_padding_attention_strided_views(ops):
    """encoder-style: q/k/v are slices of one fused buffer; key padding mask with extent."""
    B, S, H, Dh = 3, 90, 2, 128
    buf = _rand((3, B, S, H, Dh), 14, dtype=BF16)
    mask = torch.zeros(B, S, dtype=torch.uint8, device="cuda"); mask[0, :90] = 1; mask[1, :10] = 1; mask[2, :64] = 1
    out = torch.empty((B * S, H * Dh), device="cuda", dtype=BF16)
    ops.attention(buf[0], [ops.Segment(buf[1], buf[2], mask=mask, extent=ops.mask_extent(mask))], out)
    ref = _ref_attention(buf[0], [dict(k=buf[1], v=buf[2], mask=mask)], None)
    assert rel_l2(out.float(), ref) < 6e-3


@pytest.mark.parametrize("branches,rescale", [(3, False), (3, True), (1, True)])
def test_cfg_euler_tail(ops, branches, rescale):
    from oracle import echo_oracle as O
    B, S, D, lat = 2, 37, 512, 80
    P = B * S
    h = _rand((branches * P, D), 20, 2.0)
    nw, w, b = _rand((D,), 21).abs() + 0.5, _rand((lat, D), 22, D ** -0.5), _rand((lat,), 23)
    x0 = _rand((P, lat), 24)
    t, tn = 0.7243, 0.6991
    sched = torch.tensor([t, tn - t, 0, 0], device="cuda", dtype=F32)
    x = x0.clone()
    vout = torch.empty((branches * P, lat), device="cuda")
    vc = torch.empty((P, lat), device="cuda")
    ops.cfg_euler_tail(h, nw, w, b, positions=P, branches=branches, eps=1e-5, x_t=x, sched=sched, cfg_text=3.0,
                       cfg_speaker=8.0, rescale_k=1.2 if rescale else None, rescale_sigma=3.0 if rescale else None,
                       v_out=vout, v_comb_out=vc)
    hn = O.rms_norm(h.cpu(), nw.cpu(), 1e-5)
    v = hn @ w.cpu().t() + b.cpu()
    assert rel_l2(vout.cpu(), v) < 1e-5
    if branches == 3:
        a, bb, c = v.chunk(3, 0)
        v = a + 3.0 * (a - bb) + 8.0 * (a - c)
    if rescale:
        v = O.temporal_score_rescale(v, x0.cpu(), torch.tensor(t), 1.2, 3.0)
    assert rel_l2(vc.cpu(), v) < 2e-5
    tt = sched.cpu()
    assert rel_l2(x.cpu(), x0.cpu() + v * tt[1]) < 1e-5


@pytest.mark.parametrize("rescale", [False, True])
def test_cfg_combine_euler_equals_fused_tail(ops, rescale):
    """The combine-only kernel of the multi-GPU CFG-branch split = the fused tail fed with the same branch velocities."""
    B, S, D, lat = 2, 41, 256, 80
    P = B * S
    h = _rand((3 * P, D), 30, 2.0)
    nw, w, b = _rand((D,), 31).abs() + 0.5, _rand((lat, D), 32, D ** -0.5), _rand((lat,), 33)
    x0 = _rand((P, lat), 34)
    sched = torch.tensor([0.61, -0.025, 0, 0], device="cuda", dtype=F32)
    kw = dict(cfg_text=3.0, cfg_speaker=8.0, rescale_k=1.2 if rescale else None, rescale_sigma=3.0 if rescale else None)
    x_fused, v3 = x0.clone(), torch.empty((3 * P, lat), device="cuda")
    ops.cfg_euler_tail(h, nw, w, b, positions=P, branches=3, eps=1e-5, x_t=x_fused, sched=sched, v_out=v3, **kw)
    x_split, vcomb = x0.clone(), torch.empty((P, lat), device="cuda")
    ops.cfg_combine_euler(v3.view(3, P, lat), x_split, sched, v_comb_out=vcomb, **kw)
    assert rel_l2(x_split, x_fused) < 1e-6
    a, bb, c = v3.chunk(3, 0)
    if not rescale:
        assert rel_l2(vcomb, a + 3.0 * (a - bb) + 8.0 * (a - c)) < 1e-6


def test_small_kernels(ops):
    from oracle import echo_oracle as O
    t = torch.tensor([0.999, 0.5, 0.0123], device="cuda")
    assert rel_l2(ops.timestep_embedding(t, 512).cpu(), O.timestep_embedding(t.cpu(), 512)) < 1e-4
    ids = torch.tensor([0, 5, 255, 7], device="cuda", dtype=torch.int32)
    tab = _rand((256, 64), 30)
    assert torch.equal(ops.embedding(ids, tab), tab[ids.long()])
    m = torch.zeros(4, 50, dtype=torch.bool, device="cuda"); m[0, :7] = True; m[2, 49] = True; m[3, 0] = True
    assert ops.mask_extent(m).tolist() == [7, 0, 50, 1]
    x = _rand((2, 40, 128), 31)
    w, b, lw, lb = _rand((128, 7), 32, 0.4), _rand((128,), 33, 0.1), _rand((128,), 34).abs() + 0.5, _rand((128,), 35, 0.1)
    y = torch.empty((2, 40, 128), device="cuda", dtype=BF16)
    ops.dwconv_layernorm(x, w, b, lw, lb, 1e-6, y)
    xc = x.cpu().transpose(1, 2)
    ref = O.causal_conv1d(xc, w.cpu()[:, None, :], b.cpu(), groups=128).transpose(1, 2)
    ref = torch.nn.functional.layer_norm(ref, (128,), lw.cpu(), lb.cpu(), 1e-6)
    assert rel_l2(y.float().cpu(), ref) < 4e-3
    xs = _rand((1, 600, 96), 36, dtype=BF16)
    wl = _rand((7, 96), 37, 0.05)
    wav = torch.empty((1, 600), device="cuda")
    ops.conv_out_tanh(xs, wl, 0.01, wav)
    refw = torch.tanh(O.causal_conv1d(xs.float().cpu().transpose(1, 2), wl.cpu().t()[None], torch.tensor([0.01])))
    assert rel_l2(wav.cpu(), refw[:, 0]) < 1e-4


def test_flattening_point_matches_golden(ops):
    import json, os
    from conftest import GOLDEN
    idx = {c["case"]: c["index"] for c in json.load(open(os.path.join(GOLDEN, "index_golden.json")))["flatten"]}
    cases = torch.load(os.path.join(GOLDEN, "flatten_cases.pt"))
    for name, data in cases.items():
        assert int(ops.flattening_point(data.cuda().contiguous()).item()) == idx[name], name


def test_flattening_point_long_latent_vs_oracle(ops):
    """A 6400-frame latent (a long continuation; more than the 48 KB default of shared memory holds): the reference accepts
    any length (inference.py:315-330) and the index must still be the reference loop's."""
    from oracle import echo_oracle as O
    g = torch.Generator().manual_seed(31)
    for flat_from in (5000, 6399, 17):
        data = torch.cat([torch.randn(flat_from, 80, generator=g), 0.01 * torch.randn(6400 - flat_from, 80, generator=g)])
        assert int(ops.flattening_point(data.cuda().contiguous()).item()) == O.find_flattening_point(data)
