"""Is a 160-latent block of the blockwise sampler host- or device-bound?  Host time until the launches of one block have
all been issued (no sync) vs the time until the device has finished them."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from echo_tts_b200 import schema as S
from echo_tts_b200.inference import load_model_from_state_dict
from echo_tts_b200 import sampler as SP
import bench


def main():
    model = load_model_from_state_dict(S.synth_state_dict(S.dit_schema(S.ECHO_BASE, blockwise=True), init="torch"))
    ids, mask, spk, smask = [t.cuda() for t in bench.synth_inputs(1)]
    kw = dict({k: v for k, v in bench.KW.items() if k != "sequence_length"}, cfg_scale_speaker=5.0)
    real = SP.run_steps
    rec = []

    def timed_run_steps(*a, **k):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        real(*a, **k)
        t1 = time.perf_counter()
        torch.cuda.synchronize()
        rec.append(((t1 - t0) * 1e3, (time.perf_counter() - t0) * 1e3))

    SP.run_steps = timed_run_steps
    for rep in range(3):
        rec.clear()
        for _ in SP.blockwise_blocks(model, spk, smask, ids, mask, 0, [160] * 4, **kw):
            pass
    for i, (h, d) in enumerate(rec):
        print(f"block {i}: launches issued after {h:.1f} ms, device done after {d:.1f} ms")


if __name__ == "__main__":
    main()
