"""Dev experiment: per-level cost of the forward gather and sensitivity to the hashmap size."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from neuralangelo_b200 import hashgrid, _lib, build
build.build(verbose=False)
dev = torch.device("cuda:0")
n = 1 << 22
g = torch.Generator(device=dev).manual_seed(0)
p = torch.randn(n, 3, device=dev, generator=g)
p = p / p.norm(dim=-1, keepdim=True) * torch.rand(n, 1, device=dev, generator=g) ** (1 / 3)
x = ((p * 0.5 + 0.5) * 0.5 + 0.25).contiguous()
um = _lib.UnitMap(0.0, 1.0)
def timeit(fn, iters=5, warm=2):
    for _ in range(warm): fn()
    torch.cuda.synchronize(); ts = []
    for _ in range(iters):
        s, t = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); t.record(); torch.cuda.synchronize(); ts.append(s.elapsed_time(t))
    return float(np.median(ts))
out = torch.empty(1, n, 128, device=dev)
for T in (22, 21, 20, 19):
    cfg = (16, 8, T, 32, float(np.exp((np.log(2048) - np.log(32)) / 15)))
    desc = hashgrid.make_desc(*cfg)
    table = (torch.rand(desc.total_entries * 8, device=dev, generator=g) * 2 - 1) * 1e-4
    ts = [timeit(lambda: hashgrid.encode_taps(desc, table, x, um, a, ((0, 0, 0),), out=out)) for a in range(0, 17)]
    print("T", T, "cumulative ms by active levels:", [round(t, 2) for t in ts])
    print("   per-level increments:", [round(ts[i + 1] - ts[i], 3) for i in range(16)])
