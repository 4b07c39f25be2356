import torch
def t(fn, n=20):
    fn(); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(n): fn()
    g.replay(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); g.replay(); b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) * 1e3 / n
for mb, shape in ((19.7, (3, 640, 2048)), (15.7, (1920, 2048)), (5.2, (640, 2048))):
    x = torch.zeros(shape, device="cuda"); y = torch.ones(shape, device="cuda")
    print(f"{mb} MB fp32: fill_ {t(lambda: x.fill_(1.0)):.2f} us, copy_ {t(lambda: x.copy_(y)):.2f} us, add_ {t(lambda: x.add_(y)):.2f} us")
