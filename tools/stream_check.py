"""Streaming decode diagnostics on a B200: per-block deviation of DecodeStream from the one-shot decode (is the
difference uniform rounding or localised at block boundaries?) and the time per block."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from echo_tts_b200 import schema as S
from echo_tts_b200.inference import load_fish_ae_from_state_dict, ae_decode, PCAState
from oracle import echo_oracle as O


def main():
    ae = load_fish_ae_from_state_dict(S.synth_ae_state_dict(S.AE_BASE))
    comp, mean, scale = O.synth_pca()
    pca = PCAState(comp.cuda(), mean.cuda(), scale)
    z = torch.randn(1, 640, 80, device="cuda", generator=torch.Generator(device="cuda").manual_seed(9))
    full = ae_decode(ae, pca, z)
    # fp32 CPU oracle on the first 320 frames (causal: a prefix decode is the decode's prefix): both bf16 realisations
    # must sit at the same distance from it in the span where the window is full (frames 160..320)
    t0 = time.perf_counter()
    zq = ((z[:, :320].cpu() / scale) @ comp + mean).transpose(1, 2).contiguous()        # inference.py:232-235
    ref = O.decode_zq(S.synth_ae_state_dict(S.AE_BASE), S.AE_BASE, zq)
    print(f"oracle decode of 320 frames: {time.perf_counter() - t0:.1f} s on {torch.get_num_threads()} threads")
    a, b = 160 * 2048, 320 * 2048
    rl = lambda x, y: float((x - y).norm() / y.norm())
    print(f"one-shot vs oracle: frames 0..160 {rl(full[..., :a].cpu(), ref[..., :a]):.3e}, 160..320 {rl(full[..., a:b].cpu(), ref[..., a:b]):.3e}")
    st = ae.decode_stream()
    got = torch.cat([st.step_latent(z[:, i:i + 160], pca.pca_components, pca.pca_mean, pca.latent_scale) for i in (0, 160)], -1)
    print(f"stream   vs oracle: frames 0..160 {rl(got[..., :a].cpu(), ref[..., :a]):.3e}, 160..320 {rl(got[..., a:b].cpu(), ref[..., a:b]):.3e}")
    for blk in (160, 40, 8):
        st = ae.decode_stream()
        outs = []
        torch.cuda.synchronize()
        times = []
        for i in range(0, 640, blk):
            t0 = time.perf_counter()
            outs.append(st.step_latent(z[:, i:i + blk], pca.pca_components, pca.pca_mean, pca.latent_scale))
            torch.cuda.synchronize()
            times.append((time.perf_counter() - t0) * 1e3)
        got = torch.cat(outs, -1)
        errs = []
        for i in range(0, 640, 160):
            a, b = i * 2048, (i + 160) * 2048
            errs.append(float((got[..., a:b] - full[..., a:b]).norm() / full[..., a:b].norm()))
        # boundary-localised? error in the 2048 samples right after each 160-frame boundary vs elsewhere
        edge = [float((got[..., i * 2048:(i + 1) * 2048] - full[..., i * 2048:(i + 1) * 2048]).norm()
                      / full[..., i * 2048:(i + 1) * 2048].norm()) for i in (160, 320, 480)]
        print(f"block {blk}: rel-L2 per 160-frame span {['%.2e' % e for e in errs]}, first frame after a boundary "
              f"{['%.2e' % e for e in edge]}, ms/step median {sorted(times)[len(times) // 2]:.2f} (first {times[0]:.1f})")


if __name__ == "__main__":
    main()
