"""Developer microbenchmark: fused tcgen05 SDF MLP forward vs eager fp32 torch (cuBLAS)."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from neuralangelo_b200 import build
from neuralangelo_b200.mlp import MLPforNeuralSDF, get_activation
from neuralangelo_b200.fused_mlp import PackedSdfMlp
build.build(verbose=False)
dev = torch.device("cuda:0")
torch.manual_seed(0)
mlp = MLPforNeuralSDF([131, 256, 256], activ=get_activation("softplus", beta=100.0), use_weightnorm=True, geometric_init=True, out_bias=0.5).to(dev)
packed = PackedSdfMlp(mlp, 100.0)
m = 1 << 21
x = torch.randn(m, 132, device=dev)[:, 1:]
def timeit(fn, iters=5, warm=2):
    for _ in range(warm): fn()
    torch.cuda.synchronize(); ts = []
    for _ in range(iters):
        s, t = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); t.record(); torch.cuda.synchronize(); ts.append(s.elapsed_time(t))
    return float(np.median(ts))
res = {}
for name, wf in (("sdf_only", False), ("sdf_feat", True)):
    ms = timeit(lambda: packed.forward(x, with_sdf=True, with_feat=wf))
    xc = x.contiguous()
    ms_packed = timeit(lambda: packed.forward(xc, with_sdf=True, with_feat=wf))
    flop = 2 * m * (131 * 256 + 256 + (256 * 256 if wf else 0))
    res[name] = dict(ms=ms, ms_packed_rows=ms_packed, Mrows_s_packed=m / ms_packed / 1e3, Mrows_s=m / ms / 1e3, fp32_equiv_TFLOPs=flop / ms / 1e9, mma_TFLOPs=3 * 2 * m * (144 * 256 + (256 * 256 if wf else 0)) / ms / 1e9)
    with torch.no_grad():
        ms_t = timeit(lambda: mlp(x, with_sdf=True, with_feat=wf))
    res[name]["torch_fp32_ms"] = ms_t
planes = torch.randn(17, m, 8, device=dev)
res["sdf_only_planes_ms"] = timeit(lambda: packed.forward_planes(planes, m))
print(json.dumps(res, indent=1))
