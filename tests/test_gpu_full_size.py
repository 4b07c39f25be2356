"""Parity of the CUDA path AT THE BENCHMARKED SIZES (echo-tts-base architecture, 640 latents) against

  (a) fixtures written by EXECUTING the upstream reference in fp32 on CPU (oracle/gen_golden_full.py):
      BASELINE.json configs[1] (the bench workload: 40 steps, 768-padded text, 212-latent speaker reference) -- teacher-forced
      velocities at six steps + the final latent through the CUDA-graph replay path; configs[3] (B = 2, 6400-latent
      reference, Force Speaker 1.5) and configs[4] (a 160-latent block behind a 480-latent prefix) -- one guided and one
      un-guided step each;
  (b) the CPU oracle run LIVE on the box's host cores (one three-branch forward at t ~ 0.9, one single forward at t ~ 0.4).

Tolerances are BASELINE.json's: per-step velocity rel-L2 <= 1e-2 (per branch AND combined), final latent <= 2e-2.
At these sizes the engine takes paths the small fixtures never reach: 7.5 / 2.5 row tiles with a half-empty last tile, the
narrow-tail schedule, split-K projections with the deferred residual, 240- / 80-CTA attention grids with the padded text
segment skipped by extent, and 40 steps of accumulation."""
import os

import pytest
import torch

from conftest import GOLDEN, rel_l2
from echo_tts_b200 import schema as S

pytestmark = [pytest.mark.gpu, pytest.mark.slow]
V_TOL, X_TOL = 1e-2, 2e-2


@pytest.fixture(scope="module")
def base_sd():
    return S.synth_state_dict(S.dit_schema(S.ECHO_BASE, blockwise=True), init="torch")


@pytest.fixture(scope="module")
def base_model(base_sd):
    from echo_tts_b200.inference import load_model_from_state_dict
    return load_model_from_state_dict(base_sd)


def _check_steps(trace, recs, B, label):
    by_i = {st["i"]: st for st in trace}
    worst_b = worst_c = 0.0
    for rec in recs:
        st = by_i[rec["i"]]
        assert st["has_cfg"] == rec["has_cfg"], (label, rec["i"])
        n = 3 if rec["has_cfg"] else 1
        for a, b in zip(st["v_branches"].cpu().chunk(n, 0), rec["v_branches"].float().chunk(n, 0)):
            worst_b = max(worst_b, rel_l2(a, b))
        worst_c = max(worst_c, rel_l2(st["v"].cpu(), rec["v"]))
    print(f"{label}: worst per-branch velocity rel-L2 {worst_b:.3e}, worst combined velocity rel-L2 {worst_c:.3e} "
          f"(tol {V_TOL}) over steps {[r['i'] for r in recs]}")
    assert worst_b < V_TOL and worst_c < V_TOL
    return worst_b, worst_c


def test_c2_teacher_forced_vs_upstream_fixture(base_model):
    """BASELINE configs[1]: velocities of the CUDA path at the UPSTREAM sampler's own x_t (steps 0, 10, 19 guided;
    20, 30, 39 un-guided) against the upstream fp32 model's."""
    from echo_tts_b200.sampler import sample_euler_cfg_independent_guidances as sample
    g = torch.load(os.path.join(GOLDEN, "dit_base_c2.pt"))
    ids, mask, spk, smask = S.synth_request(1, seed=7)
    assert torch.equal(ids, g["ids"]) and torch.equal(spk, g["speaker_latent"])          # the bench workload's inputs
    trace = []
    sample(base_model, spk.cuda(), smask.cuda(), ids.cuda(), mask.cuda(), 0, _noise=g["noise"], _trace=trace,
           _teacher={r["i"]: r["x_in"] for r in g["steps"]}, **g["kwargs"])
    assert len(trace) == 40 and sum(st["has_cfg"] for st in trace) == 20
    for rec in g["steps"]:
        assert torch.equal(trace[rec["i"]]["x_in"].cpu().view_as(rec["x_in"]), rec["x_in"])
    _check_steps(trace, g["steps"], 1, "configs[1] teacher-forced")


def test_c2_graph_replay_final_latent_vs_upstream_fixture(base_model):
    """BASELINE configs[1]: the final latent after 40 steps through the path bench.py times (eager first call, graph
    capture, graph replay -- all three) against the upstream sampler's."""
    from echo_tts_b200.sampler import sample_euler_cfg_independent_guidances as sample
    g = torch.load(os.path.join(GOLDEN, "dit_base_c2.pt"))
    args = [t.cuda() for t in (g["speaker_latent"], g["speaker_mask"], g["ids"], g["text_mask"])]
    outs = [sample(base_model, *args, 0, _noise=g["noise"], **g["kwargs"]).cpu() for _ in range(3)]
    errs = [rel_l2(o, g["x_final"]) for o in outs]
    print(f"configs[1] final latent rel-L2 vs upstream: eager {errs[0]:.3e}, capture {errs[1]:.3e}, replay {errs[2]:.3e} "
          f"(tol {X_TOL}); upstream sampler took {g['ref_seconds']:.0f} s on CPU")
    assert torch.equal(outs[1], outs[2])
    assert max(errs) < X_TOL


def test_c2_live_oracle_forwards(base_model, base_sd):
    """One three-branch forward at t ~ 0.9 and one single forward at t ~ 0.4 at the bench shapes, the oracle computed
    here on the host cores (no fixture in between), the CUDA path through echo_dit_step."""
    from oracle import echo_oracle as O
    from echo_tts_b200.sampler import sample_euler_cfg_independent_guidances as sample
    cfg = S.ECHO_BASE
    ids, mask, spk, smask = S.synth_request(1, seed=7)
    gen = torch.Generator().manual_seed(17)
    x_hi, x_lo = torch.randn(1, 640, 80, generator=gen), 0.7 * torch.randn(1, 640, 80, generator=gen)
    kw = dict(num_steps=40, cfg_scale_text=3.0, cfg_scale_speaker=8.0, cfg_min_t=0.5, cfg_max_t=1.0, truncation_factor=None,
              rescale_k=None, rescale_sigma=None, speaker_kv_scale=None, speaker_kv_max_layers=None, speaker_kv_min_t=None,
              sequence_length=640)
    trace = []
    sample(base_model, spk.cuda(), smask.cuda(), ids.cuda(), mask.cuda(), 0, _noise=x_hi, _trace=trace,
           _teacher={4: x_hi, 24: x_lo}, **kw)
    t_sched = O.t_schedule(40)
    with torch.inference_mode():
        kv_t = O.kv_cache_text(base_sd, cfg, ids, mask)
        kv_s = O.kv_cache_speaker(base_sd, cfg, spk)
        rep3 = lambda kv: [(torch.cat([k] * 3, 0), torch.cat([v] * 3, 0)) for k, v in kv]
        tm3 = torch.cat([mask, torch.zeros_like(mask), mask], 0)
        sm3 = torch.cat([smask, smask, torch.zeros_like(smask)], 0)
        v3 = O.dit_forward(base_sd, cfg, torch.cat([x_hi] * 3, 0), torch.ones(3) * t_sched[4], tm3, sm3, rep3(kv_t), rep3(kv_s))
        v1 = O.dit_forward(base_sd, cfg, x_lo, torch.ones(1) * t_sched[24], mask, smask, kv_t, kv_s)
    vc, vt, vs = v3.chunk(3, 0)
    recs = [dict(i=4, has_cfg=True, v_branches=v3, v=vc + 3.0 * (vc - vt) + 8.0 * (vc - vs)),
            dict(i=24, has_cfg=False, v_branches=v1, v=v1)]
    assert abs(float(t_sched[4]) - 0.9) < 2e-3 and abs(float(t_sched[24]) - 0.4) < 2e-3
    _check_steps(trace, recs, 1, "configs[1] live oracle")


def test_c4_long_reference_force_speaker_vs_upstream_fixture(base_model):
    """BASELINE configs[3]: two prompts, 6400-latent references (1600 cached speaker keys, the second valid to 3000),
    Force Speaker 1.5 over 24 layers until t < 0.9: a guided step on the scaled cache and an un-guided step on the
    un-scaled one, against the upstream sampler."""
    from oracle.gen_golden_full import c4_inputs
    from echo_tts_b200.sampler import sample_euler_cfg_independent_guidances as sample
    g = torch.load(os.path.join(GOLDEN, "dit_base_c4.pt"))
    ids, mask, spk, smask, noise = c4_inputs()
    assert abs(float(spk.double().sum() + noise.double().sum() + ids.sum()) - g["input_checksum"]) < 1e-6
    dev = [t.cuda() for t in (spk, smask, ids, mask)]
    trace = []
    sample(base_model, *dev, 0, _noise=noise, _trace=trace, _teacher={1: g["steps"][1]["x_in"]}, **g["kwargs"])
    assert [st["has_cfg"] for st in trace] == [True, False]
    _check_steps(trace, g["steps"], 2, "configs[3] (B=2, 6400-latent reference, kv scale 1.5)")
    outs = [sample(base_model, *dev, 0, _noise=noise, **g["kwargs"]).cpu() for _ in range(3)]      # eager, capture, replay
    err = rel_l2(outs[2], g["x_final"])
    print(f"configs[3] latent after the 2-step schedule rel-L2 {err:.3e} (tol {X_TOL})")
    assert torch.equal(outs[1], outs[2]) and err < X_TOL and rel_l2(outs[0], g["x_final"]) < X_TOL


@pytest.mark.parametrize("incremental", [True, False])
def test_c5_block_behind_480_prefix_vs_upstream_fixture(base_model, incremental):
    """BASELINE configs[4]: the last 160-latent block behind a 480-latent prefix (120 prefix patches through the latent
    encoder, start_pos = 480 in the rotary table), prefix KV grown incrementally or recomputed, against the upstream
    blockwise sampler (inference_blockwise.py:14-123)."""
    from oracle.gen_golden_full import c5_inputs
    from echo_tts_b200.inference_blockwise import sample_blockwise_euler_cfg_independent_guidances as sample
    g = torch.load(os.path.join(GOLDEN, "dit_base_c5.pt"))
    ids, mask, spk, smask, cont, noise = c5_inputs()
    assert abs(float(cont.double().sum() + noise.double().sum() + spk.double().sum()) - g["input_checksum"]) < 1e-6
    trace = []
    out = sample(base_model, spk.cuda(), smask.cuda(), ids.cuda(), mask.cuda(), 0, continuation_latent=cont.cuda(),
                 _noises=[noise], _trace=trace, incremental_prefix=incremental, **g["kwargs"]).cpu()
    assert out.shape == (1, 640, 80) and torch.equal(out[:, :480], cont)
    for st, rec in zip(trace, g["steps"]):
        assert rel_l2(st["x_in"].cpu(), rec["x_in"]) < 1e-3
    _check_steps(trace, g["steps"], 1, f"configs[4] block (incremental_prefix={incremental})")
    err = rel_l2(out[:, 480:], g["out_block"])
    print(f"configs[4] block latent rel-L2 {err:.3e} (tol {X_TOL})")
    assert err < X_TOL
