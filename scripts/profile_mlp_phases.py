"""Developer tool: per-phase clock64 totals of the pipelined SDF-only MLP forward (CTA 0), from a -DNA_MLP_PROF build."""
import os, sys, ctypes
os.environ["NA_NVCC_EXTRA"] = "-DNA_MLP_PROF"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from neuralangelo_b200 import build, _lib
build.build(verbose=False)
from neuralangelo_b200.mlp import MLPforNeuralSDF, get_activation
from neuralangelo_b200.fused_mlp import PackedSdfMlp
dev = torch.device("cuda:0")
torch.manual_seed(0)
mlp = MLPforNeuralSDF([131, 256, 256], activ=get_activation("softplus", beta=100.0), use_weightnorm=True, geometric_init=True, out_bias=0.5).to(dev)
packed = PackedSdfMlp(mlp, 100.0)
m = 1 << 21
x = torch.randn(m, 131, device=dev)
planes = torch.randn(17, m, 8, device=dev)
L = _lib.lib()
L.na_dev_mlp_prof.argtypes = [ctypes.c_void_p, ctypes.c_int]
buf = (ctypes.c_ulonglong * 16)()
for _ in range(2):
    packed.forward_planes(planes, m)
torch.cuda.synchronize()
L.na_dev_mlp_prof(None, 1)
packed.forward_planes(planes, m)
torch.cuda.synchronize()
L.na_dev_mlp_prof(buf, 1)
tiles = (m // 128 + 147) // 148
names = ["prefetch issue", "wait d_full0", "epilogue 0", "wait a_free", "convert", "wait d_full1", "epilogue 1", "head/sdf",
         "MMA: wait a_ready", "MMA: wait d_empty0", "MMA: issue unit0", "MMA: wait d_empty1", "MMA: issue unit1"]
print("tiles per CTA", tiles)
for i, nme in enumerate(names):
    print(f"{nme:22s} {buf[i] / tiles:9.0f} cycles/tile")
print("compute total", sum(buf[i] for i in range(8)) / tiles, " MMA total", sum(buf[i] for i in range(8, 13)) / tiles)
import numpy as np
def timeit(fn, iters=5, warm=2):
    for _ in range(warm): fn()
    torch.cuda.synchronize(); ts = []
    for _ in range(iters):
        s, t = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); t.record(); torch.cuda.synchronize(); ts.append(s.elapsed_time(t))
    return float(np.median(ts))
print("planes ms (profiling build)", timeit(lambda: packed.forward_planes(planes, m)), " rows ms", timeit(lambda: packed.forward(x, with_sdf=True, with_feat=False)))
