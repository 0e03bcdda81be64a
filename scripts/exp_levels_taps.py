"""Dev experiment: per-level cost of the 7-evaluation plane-layout forward / backward on the bench workload
(cumulative over active levels; the forward still writes zeros for inactive levels, the backward skips them)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import bench
from neuralangelo_b200 import hashgrid, _lib, build
build.build(verbose=False)
dev = torch.device("cuda:0")
n = bench.N_POINTS
desc = hashgrid.make_desc(*bench.HG_CFG)
taps = bench.six_taps(bench.EPS_WORLD)
um = _lib.UnitMap(-2.0, 0.25)
g = torch.Generator(device=dev).manual_seed(1337)
table = (torch.rand(desc.total_entries * 8, device=dev, generator=g) * 2 - 1) * 1e-4
x = bench.unit_ball_points(n, 0, dev)
dout = torch.randn(len(taps), 16, n, 8, device=dev, generator=g)
out = torch.empty(len(taps), 16, n, 8, device=dev)
dtable = torch.zeros_like(table)


def timeit(fn, iters=5, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        s, t = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); t.record(); torch.cuda.synchronize(); ts.append(s.elapsed_time(t))
    return float(np.median(ts))


for name, tp in (("7 taps", taps), ("1 tap", taps[:1])):
    o, d = out[:len(tp)], dout[:len(tp)].contiguous()
    o = o.contiguous()
    f = [timeit(lambda: hashgrid.encode_taps_planes(desc, table, x, um, a, tp, out=o)) for a in range(0, 17)]
    b = [timeit(lambda: hashgrid.scatter_taps_planes(desc, x, um, a, tp, d, dtable)) for a in range(0, 17)]
    print(name, "fwd cumulative:", [round(t, 2) for t in f])
    print(name, "fwd per level :", [round(f[i + 1] - f[i], 3) for i in range(16)])
    print(name, "bwd cumulative:", [round(t, 2) for t in b])
    print(name, "bwd per level :", [round(b[i + 1] - b[i], 3) for i in range(16)])


def morton_key(xu, bits):
    q = (xu.clamp(0, 1 - 1e-7) * (1 << bits)).long()
    key = torch.zeros(xu.shape[0], dtype=torch.long, device=xu.device)
    for b in range(bits):
        for d in range(3):
            key |= ((q[:, d] >> b) & 1) << (3 * b + d)
    return key


xu = (x + 2.0) * 0.25
for bits in (5, 7, 10):
    perm = torch.argsort(morton_key(xu, bits))
    xs = x[perm].contiguous()
    for name, tp in (("7 taps", taps), ("1 tap", taps[:1])):
        o, d = out[:len(tp)].contiguous(), dout[:len(tp)].contiguous()
        f = [timeit(lambda: hashgrid.encode_taps_planes(desc, table, xs, um, a, tp, out=o)) for a in range(0, 17)]
        b = [timeit(lambda: hashgrid.scatter_taps_planes(desc, xs, um, a, tp, d, dtable)) for a in range(0, 17)]
        print(f"sorted morton {bits} bits/axis,", name, "fwd total", round(f[16], 2), "per level:", [round(f[i + 1] - f[i], 3) for i in range(16)])
        print(f"sorted morton {bits} bits/axis,", name, "bwd total", round(b[16], 2), "per level:", [round(b[i + 1] - b[i], 3) for i in range(16)])
