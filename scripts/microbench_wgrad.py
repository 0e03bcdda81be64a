"""Developer microbenchmark: na_linear_wgrad against the library form (split-K bmm + blocked column sum) on the
weight-gradient shapes of one default-config training step."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from neuralangelo_b200 import build
from neuralangelo_b200.mlp import linear_wgrad, tn_matmul, colsum
build.build(verbose=False)
dev = torch.device("cuda:0")
torch.backends.cuda.matmul.allow_tf32 = True
shapes = [("sdf taps l0", 262144, 256, 131), ("sdf l0", 65536, 256, 131), ("sdf l1", 65536, 256, 256), ("sdf head", 327680, 1, 256),
          ("rgb l0", 65536, 256, 278), ("rgb l1-3", 65536, 256, 256), ("rgb head", 65536, 3, 256),
          ("bg l0", 16384, 256, 84), ("bg l1-6", 16384, 256, 256), ("bg skip", 16384, 256, 340), ("bg l7", 16384, 257, 256),
          ("bg rgb0", 16384, 128, 272), ("bg rgb head", 16384, 3, 128)]


def timed(fn, iters=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return float(np.median(ts)) * 1e3


tot_o = tot_l = 0.0
for name, m, n, k in shapes:
    dy = torch.randn(m, n, device=dev) * 1e-6
    x = torch.relu(torch.randn(m, k, device=dev))
    ours = timed(lambda: linear_wgrad(dy, x))
    lib = timed(lambda: (tn_matmul(dy, x), colsum(dy)))
    by = m * (n + k) * 4
    print(f"{name:12s} M={m:6d} N={n:3d} K={k:3d}  ours {ours:7.1f} us ({by / ours / 1e6:6.2f} TB/s)   library TF32 {lib:7.1f} us")
    tot_o += ours; tot_l += lib
print(f"sum over the listed shapes: ours {tot_o:.0f} us, library {tot_l:.0f} us")
