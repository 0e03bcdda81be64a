"""Dev experiment: ways to take column sums (bias gradients) of tall [M,256] fp32 matrices."""
import numpy as np, torch
dev = torch.device("cuda:0")
def timeit(fn, iters=10, warm=3):
    for _ in range(warm): fn()
    torch.cuda.synchronize(); ts = []
    for _ in range(iters):
        s, t = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); t.record(); torch.cuda.synchronize(); ts.append(s.elapsed_time(t))
    return float(np.median(ts)) * 1e3
torch.backends.cuda.matmul.allow_tf32 = True
for m in (262144, 65536, 16384):
    x = torch.randn(m, 256, device=dev)
    ones = torch.ones(m, device=dev)
    ones_row = torch.ones(1, m, device=dev)
    ref = x.double().sum(0)
    for name, fn in (("sum(0)", lambda: x.sum(0)),
                     ("ones@x", lambda: (ones_row @ x)[0]),
                     ("mv", lambda: torch.mv(x.t(), ones)),
                     ("view64.sum(1).sum(0)", lambda: x.view(64, -1, 256).sum(1).sum(0)),
                     ("view512.sum(1).sum(0)", lambda: x.view(512, -1, 256).sum(1).sum(0))):
        err = (fn().double() - ref).abs().max().item()
        print(m, f"{name:24s} {timeit(fn):8.1f} us   err {err:.2e}   ({m*1024/1e3/timeit(fn):.0f} GB/s)")
