"""Developer microbenchmark for the hash-grid kernels (not bench.py): CUDA-event timings."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from neuralangelo_b200 import hashgrid, _lib, build

build.build(verbose=False)
dev = torch.device("cuda:0")
cfg = (16, 8, 22, 32, float(np.exp((np.log(2048) - np.log(32)) / 15)))
desc = hashgrid.make_desc(*cfg)
n = 1 << 22
g = torch.Generator(device=dev).manual_seed(0)
p = torch.randn(n, 3, device=dev, generator=g)
p = p / p.norm(dim=-1, keepdim=True) * torch.rand(n, 1, device=dev, generator=g) ** (1 / 3)
x = ((p * 0.5 + 0.5) * 0.5 + 0.25).contiguous()
table = (torch.rand(desc.total_entries * 8, device=dev, generator=g) * 2 - 1) * 1e-4
um = _lib.UnitMap(0.0, 1.0)
e = float(np.float32(1.0 / 8192))
six = ((0, 0, 0), (e, 0, 0), (-e, 0, 0), (0, e, 0), (0, -e, 0), (0, 0, e), (0, 0, -e))


ITERS=int(os.environ.get('NA_MB_ITERS','5')); WARM=int(os.environ.get('NA_MB_WARM','2'))
def timeit(fn, iters=None, warm=None):
    iters = ITERS if iters is None else iters; warm = WARM if warm is None else warm
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        s, t = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); t.record(); torch.cuda.synchronize()
        ts.append(s.elapsed_time(t))
    return float(np.median(ts)), float(np.min(ts))

res = {}
for name, taps in (("fwd1", six[:1]), ("fwd7", six)):
    out = torch.empty(len(taps), n, 128, device=dev)
    med, mn = timeit(lambda: hashgrid.encode_taps(desc, table, x, um, 16, taps, out=out))
    by = len(taps) * n * (12 + 512) + table.numel() * 4
    res[name] = dict(ms=med, min_ms=mn, algo_GB=by / 1e9, GBps=by / med / 1e6)
    dout = torch.randn_like(out)
    dt = torch.zeros_like(table)
    med, mn = timeit(lambda: hashgrid.scatter_taps(desc, x, um, 16, taps, dout, dt))
    by = len(taps) * n * (12 + 512) + 2 * table.numel() * 4
    res[name.replace("fwd", "bwd")] = dict(ms=med, min_ms=mn, algo_GB=by / 1e9, GBps=by / med / 1e6)
    del out, dout, dt
# level-major plane layout
for name, taps in (("fwd1_planes", six[:1]), ("fwd7_planes", six)):
    out = torch.empty(len(taps), 16, n, 8, device=dev)
    med, mn = timeit(lambda: hashgrid.encode_taps_planes(desc, table, x, um, 16, taps, out=out))
    by = len(taps) * n * (12 + 512) + table.numel() * 4
    res[name] = dict(ms=med, min_ms=mn, algo_GB=by / 1e9, GBps=by / med / 1e6)
    dout = torch.randn_like(out)
    dt = torch.zeros_like(table)
    med, mn = timeit(lambda: hashgrid.scatter_taps_planes(desc, x, um, 16, taps, dout, dt))
    by = len(taps) * n * (12 + 512) + 2 * table.numel() * 4
    res[name.replace("fwd", "bwd")] = dict(ms=med, min_ms=mn, algo_GB=by / 1e9, GBps=by / med / 1e6)
    del out, dout, dt
# sorted points (spatially coherent) for comparison
xs = x[torch.argsort((x * 64).floor().mul(torch.tensor([4096., 64., 1.], device=dev)).sum(-1))].contiguous()
out = torch.empty(1, n, 128, device=dev)
med, mn = timeit(lambda: hashgrid.encode_taps(desc, table, xs, um, 16, six[:1], out=out))
res["fwd1_sorted"] = dict(ms=med, min_ms=mn)
print(json.dumps(res, indent=1))
