import sys, torch
sys.path.insert(0, "/root/repo")
from echo_tts_b200 import ops
T, Cc = 1310720, 96
g = torch.Generator(device="cuda").manual_seed(0)
x = torch.randn((1, T, Cc), device="cuda", generator=g).to(torch.bfloat16)
w = torch.randn((7, Cc), device="cuda", generator=g) * 0.1
wav = torch.empty((1, T), device="cuda")
for _ in range(3): ops.conv_out_tanh(x, w, 0.01, wav)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20): ops.conv_out_tanh(x, w, 0.01, wav)
e1.record(); torch.cuda.synchronize()
us = e0.elapsed_time(e1) * 1e3 / 20
xf = torch.nn.functional.pad(x.float()[0].t()[None], (6, 0))
ref = torch.tanh(torch.nn.functional.conv1d(xf, w.t()[None].contiguous()) + 0.01)[0, 0]
err = float((wav[0] - ref).abs().max())
print("conv_out_tanh T=%d C=%d: %.1f us = %.0f GB/s (of ~6550), max abs err %.2e" % (T, Cc, us, (T * Cc * 2 + T * 4) / us / 1e3, err))
